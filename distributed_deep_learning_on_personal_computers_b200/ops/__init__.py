"""Fused ops: sm_100a kernels (csrc/) with plain-PyTorch oracles (ops.reference)."""
from . import loader, reference, functional  # noqa: F401
