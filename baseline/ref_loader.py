"""Loads the reference's OWN model classes from its unmodified script, without importing it.

The reference is a single script whose module level cannot be imported on these boxes: it looks
the host name up in a hard-coded table (``R:248-249`` -> ``KeyError``), imports ``mgzip``,
``imageio``, ``numba`` and ``matplotlib`` (absent), and opens sockets.  The model section
(``R:570-656``) is self-contained, so we read the file, cut that section out by its own markers
and ``exec`` it in a fresh namespace with the global ``NN_in_model`` (``R:687``) set by the caller.
Nothing is vendored: the script is looked up under ``baseline/_ref/`` (git-ignored copy that
travels to the GPU box) or ``/root/reference`` at run time.
"""
from __future__ import annotations

import os
from types import SimpleNamespace

_HERE = os.path.dirname(os.path.abspath(__file__))
_NAME = "Vaihingen PyTorch 2 (кластер).py"
CANDIDATES = [os.path.join(_HERE, "_ref", _NAME), os.path.join("/root/reference", _NAME)]


def find_script() -> str:
    for p in CANDIDATES:
        if os.path.exists(p):
            return p
    raise FileNotFoundError("reference script not found in " + " or ".join(CANDIDATES))


def load_reference_models(nn_in_model: int = 2, script: str | None = None) -> SimpleNamespace:
    """Returns a namespace with the reference's DoubleConv, DownBlock, UpBlock, UNet classes."""
    path = script or find_script()
    with open(path, encoding="utf-8") as fh:
        lines = fh.read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith("class DoubleConv"))
    end = next(i for i, l in enumerate(lines) if i > start and l.startswith("import imageio"))
    src = "import torch\nimport torch.nn as nn\nimport torch.nn.functional as F\n" + \
        "\n".join(lines[start:end])
    ns = {"NN_in_model": nn_in_model, "__name__": "reference_models"}
    exec(compile(src, path, "exec"), ns)
    return SimpleNamespace(DoubleConv=ns["DoubleConv"], DownBlock=ns["DownBlock"],
                           UpBlock=ns["UpBlock"], UNet=ns["UNet"], namespace=ns, path=path)


def reference_unet(base_channels: int = 32, out_classes: int = 6, mode: str = "conv_transpose"):
    """The reference's UNet at a given width.  Its width knob is the integer divisor
    ``NN_in_model`` of 64 (R:625-641); widths above 64 use a fractional divisor, which the same
    ``64//N`` expressions accept (``64 // 0.5 == 128.0`` -> cast by a thin int wrapper)."""
    if 64 % base_channels == 0:
        return load_reference_models(64 // base_channels).UNet(out_classes, mode)
    if base_channels % 64:
        raise ValueError("base_channels must divide 64 or be a multiple of 64")

    class _Div:                      # makes `64 // N`-style expressions yield ints for N < 1
        def __init__(self, mult):
            self.mult = mult

        def __rfloordiv__(self, other):
            return int(other * self.mult)

    return load_reference_models(_Div(base_channels // 64)).UNet(out_classes, mode)
