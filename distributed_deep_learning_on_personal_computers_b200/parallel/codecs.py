"""Gradient wire codecs (the lossy stage of the reference's gradient exchange).

The reference switches on a global ``model_bytes`` (``R:25``):

* ``'float16'``: every gradient is divided by ONE global ``max|g|`` taken over all parameters
  (``R:451-463``), multiplied by 100, rounded, and stored as fp16 (``R:484-487``) -- at most
  201 distinct levels.  Decoding is ``q / 100 * max`` (``R:543``).  Here: codec ``q201``.
* ``'int8'``: the same with a factor of 10 and int8 storage (``R:474``, ``R:533``): 21 levels.
  Here: codec ``q21``.
* ``'float32'``: intended pass-through, broken in the reference. Here: ``fp32``.

We keep those three for capability parity and add the ones that are useful on NVLink: honest
``bf16`` / ``fp16`` casts and max-abs scaled ``int8s`` (255 levels) / ``fp8`` (e4m3).  A codec works on a flat
fp32 tensor (the gradient bucket storage) and is a pure function pair::

    payload, scale = codec.encode(flat)      # scale is a 0-dim fp32 tensor (1.0 if unused)
    flat_hat       = codec.decode(payload, scale)

A zero gradient (``max == 0``) crashes the reference (Q4); here it encodes to zeros.
On a CUDA tensor with the extension loaded the work is done by ``csrc/elem/codec.cu`` (one amax
launch over the whole flat buffer, one quantize launch); the torch code below is the oracle and
the CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

_FP8 = getattr(torch, "float8_e4m3fn", None)


@dataclass(frozen=True)
class Codec:
    name: str
    wire_dtype: torch.dtype
    levels: Optional[float]      # multiplier S in round(g / max * S); None = plain cast
    needs_scale: bool

    @property
    def wire_bytes(self) -> int:
        return torch.empty((), dtype=self.wire_dtype).element_size()

    # -- oracle implementation -----------------------------------------------------------
    def amax(self, flat: torch.Tensor) -> torch.Tensor:
        if not self.needs_scale or flat.numel() == 0:
            return torch.ones((), dtype=torch.float32, device=flat.device)
        return flat.detach().abs().max().to(torch.float32)

    def encode(self, flat: torch.Tensor, scale: Optional[torch.Tensor] = None
               ) -> Tuple[torch.Tensor, torch.Tensor]:
        if scale is None:
            scale = self.amax(flat)
        if flat.is_cuda:
            from ..ops import functional as Fn
            if Fn.native_available():
                return Fn.codec_encode(flat, scale, self), scale
        return _encode_torch(self, flat, scale), scale

    def decode(self, payload: torch.Tensor, scale: torch.Tensor,
               out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
        if payload.is_cuda:
            from ..ops import functional as Fn
            if Fn.native_available():
                return Fn.codec_decode(payload, scale, self, out=out, accumulate=accumulate)
        val = _decode_torch(self, payload, scale)
        if out is None:
            return val
        if accumulate:
            out.add_(val)
        else:
            out.copy_(val)
        return out

    def max_abs_error(self, scale: float) -> float:
        """Worst-case round-trip error for values inside ``[-scale, scale]``."""
        if self.levels is not None:
            return scale / (2.0 * self.levels)
        if self.wire_dtype == torch.bfloat16:
            return scale * 2.0 ** -8
        if self.wire_dtype == torch.float16:
            # half an ulp of a normal number, or half the subnormal spacing (2^-24) below 2^-14
            return max(scale * 2.0 ** -11, 2.0 ** -25)
        if _FP8 is not None and self.wire_dtype == _FP8:
            return scale * 2.0 ** -4
        return 0.0


def _encode_torch(c: Codec, flat: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    g = flat.detach().to(torch.float32)
    if c.levels is not None:
        # Same arithmetic order as R:487 / R:474: (g / max) * S, then round, then cast.
        safe = torch.where(scale > 0, scale, torch.ones_like(scale))
        q = torch.round(g / safe * c.levels)
        q = torch.where(scale > 0, q, torch.zeros_like(q))          # Q4 guard
        return q.to(c.wire_dtype)
    if c.name == "fp8":
        safe = torch.where(scale > 0, scale, torch.ones_like(scale))
        return (g / safe * 448.0).clamp_(-448.0, 448.0).to(c.wire_dtype)
    return g.to(c.wire_dtype)


def _decode_torch(c: Codec, payload: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    q = payload.to(torch.float32)
    if c.levels is not None:
        return q / c.levels * scale.to(q.device)                     # R:543 / R:533
    if c.name == "fp8":
        return q / 448.0 * scale.to(q.device)
    return q


_REGISTRY: Dict[str, Codec] = {
    "fp32": Codec("fp32", torch.float32, None, False),
    "bf16": Codec("bf16", torch.bfloat16, None, False),
    "fp16": Codec("fp16", torch.float16, None, False),
    "q201": Codec("q201", torch.float16, 100.0, True),
    "q21": Codec("q21", torch.int8, 10.0, True),
    "int8s": Codec("int8s", torch.int8, 127.0, True),
}
if _FP8 is not None:
    _REGISTRY["fp8"] = Codec("fp8", _FP8, None, True)


def get_codec(name: str) -> Codec:
    from ..config import REFERENCE_CODEC_ALIASES
    name = REFERENCE_CODEC_ALIASES.get(name, name)
    try:
        return _REGISTRY[name]
    except KeyError:
        raise ValueError(f"unknown codec {name!r}; have {sorted(_REGISTRY)}") from None


def reference_restatement(g: torch.Tensor, mode: str) -> torch.Tensor:
    """Five-line restatement of the reference's encode->decode round trip (R:487+R:543 for
    ``'float16'``, R:474+R:533 for ``'int8'``); used by the tests as the parity oracle."""
    mx = g.abs().max()
    if mode == "float16":
        q = torch.round(g / mx * 100).to(torch.float16)
        return q.to(torch.float32) / 100 * mx
    q = torch.round(g / mx * 10).to(torch.int8)
    return q.to(torch.float32) / 10 * mx
