"""Checkpoint / resume.

The reference never saves anything (``torch.save`` only appears in comments, ``R:356``, ``R:393``);
BASELINE.json nevertheless fixes the layout: a torch ``state_dict`` with the reference ``UNet``'s
keys.  A checkpoint is one ``torch.save`` of::

    {"model": OrderedDict (166 entries at the shipped width, plain contiguous CPU tensors),
     "optimizer": torch.optim-style state dict, "epoch": int, "step": int, "config": dict}

Rank 0 writes (atomically: temp file + rename); every rank loads.  ``model`` loads directly into
the reference's ``UNet().load_state_dict`` and the other way round.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch


def save(path: str, model, optimizer, epoch: int, step: int, config: Dict,
         extra: Optional[Dict] = None) -> str:
    payload = {"model": model.portable_state_dict() if hasattr(model, "portable_state_dict")
               else {k: v.detach().cpu() for k, v in model.state_dict().items()},
               "optimizer": optimizer.state_dict() if optimizer is not None else None,
               "epoch": int(epoch), "step": int(step), "config": dict(config)}
    if extra:
        payload.update(extra)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    tmp = path + ".tmp"
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return path


def latest(out_dir: str, names=("checkpoint.pt", "abort.pt")) -> Optional[str]:
    """The most recently written checkpoint of a run directory (the per-epoch one or the one rank 0 leaves when a
    peer is lost), or None."""
    found = [os.path.join(out_dir, n) for n in names if os.path.isfile(os.path.join(out_dir, n))]
    return max(found, key=os.path.getmtime) if found else None


def load(path: str, model, optimizer=None, map_location="cpu", trusted: bool = False) -> Dict:
    """``trusted=False`` (default) uses torch's restricted unpickler: a checkpoint holds only tensors, dicts,
    lists, tuples and scalars, so nothing else needs to be constructible from the file (the star transport
    avoids pickle for the same reason).  ``trusted=True`` (or ``B200SEG_TRUSTED_CHECKPOINT=1``) falls back
    to the full unpickler for files written by other tools."""
    trusted = trusted or os.environ.get("B200SEG_TRUSTED_CHECKPOINT") == "1"
    ckpt = torch.load(path, map_location=map_location, weights_only=not trusted)
    with torch.no_grad():
        own = model.state_dict()
        missing = [k for k in own if k not in ckpt["model"]]
        unexpected = [k for k in ckpt["model"] if k not in own]
        if missing or unexpected:
            raise KeyError(f"checkpoint mismatch: missing={missing[:4]} unexpected={unexpected[:4]}")
        for k, v in ckpt["model"].items():
            own[k].copy_(v)              # copy_ honours the flat-buffer strides
    if optimizer is not None and ckpt.get("optimizer") is not None:
        optimizer.load_state_dict(ckpt["optimizer"])
    return ckpt
