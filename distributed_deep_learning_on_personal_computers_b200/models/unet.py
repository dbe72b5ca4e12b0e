"""U-Net for tile segmentation, with the reference's ``state_dict`` layout.

Parity contract: the module tree below produces exactly the 166
``state_dict`` entries of the reference's ``UNet(out_classes=6)`` (``R:620-656``; blocks
``R:575-617``), in the same ``parameters()`` order, so a checkpoint written here loads into the
reference class and vice versa.  What differs is everything underneath:

* the width is a ``base_channels`` value (32 = the shipped ``64 // NN_in_model``, ``R:625-641``;
  64 and 128 are the BASELINE.json widths) instead of a global divisor;
* all parameters can be re-homed into ONE flat fp32 buffer (``flatten_parameters``) with a
  matching flat gradient buffer, conv weights stored ``[Cout][kh][kw][Cin]`` (= torch
  ``channels_last`` strides of the logical ``(Cout, Cin, kh, kw)`` tensor).  The flat buffers are
  what the sm_100a kernels, the fused optimizer and the gradient all-reduce work on;
* ``forward`` here is the plain-PyTorch oracle / CPU path.  On a B200 the training step is run
  by ``models.engine.NativeEngine`` over the same parameters.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch
import torch.nn as nn

FLAT_ALIGN = 64  # elements; keeps every tensor 256 B (fp32) / 128 B (bf16) aligned for TMA


def _conv_bn_relu_pair(cin: int, cout: int) -> nn.Sequential:
    # Sequential indices 0 conv, 1 BN, 2 ReLU, 3 conv, 4 BN, 5 ReLU (R:578-585).
    layers = []
    for c_in in (cin, cout):
        layers += [nn.Conv2d(c_in, cout, kernel_size=3, padding=1),
                   nn.BatchNorm2d(cout),
                   nn.ReLU(inplace=True)]
    return nn.Sequential(*layers)


class DoubleConv(nn.Module):
    """conv3x3-BN-ReLU twice (R:575-588). The attribute is called ``double_conv`` on purpose."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.double_conv = _conv_bn_relu_pair(cin, cout)

    def forward(self, x):
        return self.double_conv(x)


class DownBlock(nn.Module):
    """DoubleConv then 2x2 max-pool; returns (pooled, skip) (R:591-600)."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.double_conv = DoubleConv(cin, cout)
        self.down_sample = nn.MaxPool2d(2)

    def forward(self, x):
        skip = self.double_conv(x)
        return self.down_sample(skip), skip


class UpBlock(nn.Module):
    """x2 up-sampling, concat ``[up, skip]`` on channels, DoubleConv (R:603-617)."""

    def __init__(self, c_up: int, c_skip: int, cout: int, mode: str):
        super().__init__()
        if mode == "conv_transpose":
            self.up_sample = nn.ConvTranspose2d(c_up, c_up, kernel_size=2, stride=2)
        elif mode == "bilinear":
            self.up_sample = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
        else:
            raise ValueError("Unsupported `up_sample_mode` (conv_transpose | bilinear)")
        self.double_conv = DoubleConv(c_up + c_skip, cout)

    def forward(self, below, skip):
        return self.double_conv(torch.cat([self.up_sample(below), skip], dim=1))


def encoder_widths(b: int) -> List[int]:
    """Output channels of down_conv1..5 (R:625-629): b, 2b, 4b, 8b, 8b."""
    return [b, 2 * b, 4 * b, 8 * b, 8 * b]


@dataclass
class ConvSpec:
    """One 3x3 conv+BN+ReLU stage as the native engine sees it."""
    name: str          # state_dict prefix of the conv, e.g. "down_conv1.double_conv.double_conv.0"
    bn: str            # prefix of the BatchNorm that follows
    cin: int
    cout: int
    level: int         # spatial level: resolution is tile >> level


class UNet(nn.Module):
    """Reference-compatible U-Net (R:620-656)."""

    def __init__(self, out_classes: int = 2, up_sample_mode: str = "conv_transpose",
                 base_channels: int = 32, in_channels: int = 3):
        super().__init__()
        self.up_sample_mode = up_sample_mode
        self.base_channels = b = int(base_channels)
        self.in_channels = in_channels
        self.out_classes = out_classes
        w = encoder_widths(b)
        cin = in_channels
        for i, cout in enumerate(w, start=1):
            setattr(self, f"down_conv{i}", DownBlock(cin, cout))
            cin = cout
        self.double_conv = DoubleConv(w[4], w[4])                      # bottleneck (R:631)
        # decoder: (channels coming up, skip channels, output channels)  (R:633-637)
        dec = {5: (w[4], w[4], w[3]), 4: (w[3], w[3], w[3]), 3: (w[3], w[2], w[2]),
               2: (w[2], w[1], w[1]), 1: (w[1], w[0], w[0])}
        for i in (5, 4, 3, 2, 1):
            c_up, c_skip, cout = dec[i]
            setattr(self, f"up_conv{i}", UpBlock(c_up, c_skip, cout, up_sample_mode))
        self.conv_last = nn.Conv2d(w[0], out_classes, kernel_size=1)
        self._flat: Dict[str, torch.Tensor] = {}

    # ---- oracle / CPU forward ------------------------------------------------------------
    def forward(self, x):
        skips = []
        for i in range(1, 6):
            x, s = getattr(self, f"down_conv{i}")(x)
            skips.append(s)
        x = self.double_conv(x)
        for i in (5, 4, 3, 2, 1):
            x = getattr(self, f"up_conv{i}")(x, skips[i - 1])
        return self.conv_last(x)

    # ---- structure description for the native engine ---------------------------------------
    def conv_specs(self) -> "OrderedDict[str, Tuple[ConvSpec, ConvSpec]]":
        out: "OrderedDict[str, Tuple[ConvSpec, ConvSpec]]" = OrderedDict()

        def pair(prefix: str, level: int):
            seq = self.get_submodule(prefix)
            c0, c1 = seq[0], seq[3]
            return (ConvSpec(f"{prefix}.0", f"{prefix}.1", c0.in_channels, c0.out_channels, level),
                    ConvSpec(f"{prefix}.3", f"{prefix}.4", c1.in_channels, c1.out_channels, level))

        for i in range(1, 6):
            out[f"down_conv{i}"] = pair(f"down_conv{i}.double_conv.double_conv", i - 1)
        out["double_conv"] = pair("double_conv.double_conv", 5)
        for i in (5, 4, 3, 2, 1):
            out[f"up_conv{i}"] = pair(f"up_conv{i}.double_conv.double_conv", i - 1)
        return out

    # ---- flat parameter storage ----------------------------------------------------------
    @staticmethod
    def _internal_view(flat: torch.Tensor, off: int, p: torch.Tensor, is_convT: bool) -> torch.Tensor:
        """Logical-shape view of ``flat[off:off+numel]`` in the kernels' memory order."""
        n = p.numel()
        seg = flat[off:off + n]
        if p.dim() != 4:
            return seg.view(p.shape)
        if is_convT:
            # logical (Cin, Cout, 2, 2); memory [dy][dx][Cout][Cin]  (GEMM B operand, K = Cin)
            ci, co, kh, kw = p.shape
            return seg.view(kh, kw, co, ci).permute(3, 2, 0, 1)
        co, ci, kh, kw = p.shape
        return seg.view(co, kh, kw, ci).permute(0, 3, 1, 2)     # == channels_last strides

    def flatten_parameters(self, device=None) -> Dict[str, torch.Tensor]:
        """Move every parameter into one flat fp32 buffer (+ a zeroed flat ``.grad`` buffer).

        Layout: first the *bulk* tensors — the 3x3 / transposed-conv weights whose gradients the
        tensor-core wgrad kernels write (everything but the 3-channel stem and the 1x1 head) — in
        ``parameters()`` order (the reference's wire order, R:453), then all the small tensors (BN
        affine, biases, stem, head).  The split lets the fused exchange treat the two regions
        differently (``parallel/allreduce.py``).  Returns ``{"param", "grad", "offsets", "bulk_elems", "bulk_names"}``;
        ``offsets[name] = (start, numel)`` in elements.  Nothing depends on the order: every consumer
        goes through ``offsets`` and checkpoints are by name.
        """
        params = list(self.named_parameters())
        device = device or params[0][1].device
        convT = {n for n, m in self.named_modules() if isinstance(m, nn.ConvTranspose2d)}

        def is_bulk(name, p):
            if p.dim() != 4 or p.shape[-1] * p.shape[-2] == 1:
                return False
            cin = p.shape[0] if name.rsplit(".", 1)[0] in convT else p.shape[1]
            return cin % 8 == 0

        offsets: "OrderedDict[str, Tuple[int, int]]" = OrderedDict()
        total = 0
        bulk_elems = 0
        bulk_names = set()
        for want_bulk in (True, False):
            for name, p in params:
                if is_bulk(name, p) != want_bulk:
                    continue
                if want_bulk:
                    bulk_names.add(name)
                offsets[name] = (total, p.numel())
                total += (p.numel() + FLAT_ALIGN - 1) // FLAT_ALIGN * FLAT_ALIGN
            if want_bulk:
                bulk_elems = total
        flat_p = torch.zeros(total, dtype=torch.float32, device=device)
        flat_g = torch.zeros(total, dtype=torch.float32, device=device)
        with torch.no_grad():
            for name, p in params:
                off, _ = offsets[name]
                is_t = name.rsplit(".", 1)[0] in convT and p.dim() == 4
                v = self._internal_view(flat_p, off, p, is_t)
                v.copy_(p.detach().to(device=device, dtype=torch.float32))
                p.data = v
                p.grad = self._internal_view(flat_g, off, p, is_t)
        for _, buf in self.named_buffers():
            buf.data = buf.data.to(device)
        self._flat = {"param": flat_p, "grad": flat_g, "offsets": offsets, "bulk_elems": bulk_elems,
                      "bulk_names": bulk_names}
        return self._flat

    @property
    def flat(self) -> Dict[str, torch.Tensor]:
        if not self._flat:
            raise RuntimeError("call flatten_parameters() first")
        return self._flat

    def portable_state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        """``state_dict`` with plain contiguous CPU tensors (the checkpoint layout, Appendix A)."""
        return OrderedDict((k, v.detach().cpu().contiguous().clone())
                           for k, v in self.state_dict().items())


def count_parameters(model: nn.Module) -> int:
    return sum(p.numel() for p in model.parameters())


def conv_flops_per_tile(base_channels: int, tile: int, classes: int = 6, in_channels: int = 3) -> float:
    """Forward conv FLOPs (2*MACs) for one tile; 113.4 GFLOP at b=32, 512x512."""
    b = base_channels
    w = encoder_widths(b)
    fl = 0.0
    cin = in_channels
    for lvl, cout in enumerate(w):
        px = (tile >> lvl) ** 2
        fl += 2.0 * px * 9 * (cin * cout + cout * cout)
        cin = cout
    px = (tile >> 5) ** 2
    fl += 2.0 * px * 9 * 2 * w[4] * w[4]
    dec = {5: (w[4], w[4], w[3]), 4: (w[3], w[3], w[3]), 3: (w[3], w[2], w[2]),
           2: (w[2], w[1], w[1]), 1: (w[1], w[0], w[0])}
    for i in (5, 4, 3, 2, 1):
        c_up, c_skip, cout = dec[i]
        px_lo = (tile >> i) ** 2
        px_hi = (tile >> (i - 1)) ** 2
        fl += 2.0 * px_lo * 4 * c_up * c_up                       # convT 2x2
        fl += 2.0 * px_hi * 9 * ((c_up + c_skip) * cout + cout * cout)
    fl += 2.0 * tile * tile * w[0] * classes
    return fl
