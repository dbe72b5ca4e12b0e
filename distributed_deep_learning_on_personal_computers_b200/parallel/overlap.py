"""Gradient buckets for the exchange that runs under the backward pass.

The reference sends ONE whole-model message after backward (``R:756-763``; SURVEY 0.2: "no DDP, no
buckets, no overlap").  Here the flat gradient buffer (``models.unet.UNet.flatten_parameters``) is cut
into buckets in the order in which backward *completes* them, so that every bucket but the last can be
reduced over NVLink while the remaining layers still compute (``csrc/comm/overlap.cu``).

``plan_buckets`` is pure bookkeeping (CPU-testable):

* ``order`` lists the bulk parameters (3x3 / transposed-conv weights, written by the tensor-core wgrad
  kernels) in backward completion order; each belongs to a *block* (``up_conv1`` ... ``down_conv1``).
* blocks are appended to the current bucket until it holds ``bucket_mb`` MiB; the LAST block always goes
  into the final bucket together with everything that is only complete when backward ends (BN affine
  gradients, biases, the stem and the head: the tail of the flat buffer).
* a bucket is a few contiguous ranges of the flat buffer (blocks are contiguous in it) plus the name of
  the parameter whose weight-gradient launch completes it: that launch carries the all-reduce trigger.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

MAX_RANGES = 4


@dataclass
class Bucket:
    index: int
    ranges: List[Tuple[int, int]] = field(default_factory=list)   # (offset, count) in elements
    params: List[str] = field(default_factory=list)
    trigger: Optional[str] = None        # parameter whose wgrad launch completes the bucket (None: final)
    final: bool = False

    @property
    def elems(self) -> int:
        return sum(n for _, n in self.ranges)


def _merge(ranges: List[Tuple[int, int]]) -> List[Tuple[int, int]]:
    out: List[Tuple[int, int]] = []
    for off, n in sorted(ranges):
        if out and out[-1][0] + out[-1][1] == off:
            out[-1] = (out[-1][0], out[-1][1] + n)
        else:
            out.append((off, n))
    return out


def block_of(name: str) -> str:
    return name.split(".", 1)[0]


def plan_buckets(offsets: Dict[str, Tuple[int, int]], order: Sequence[str], total_elems: int,
                 bucket_mb: float = 8.0, align: int = 64, max_buckets: int = 32) -> List[Bucket]:
    """``offsets``: name -> (start, numel) of every parameter in the flat buffer; ``order``: the bulk
    parameters in backward completion order; ``total_elems``: length of the flat buffer (padded)."""
    def padded(n: int) -> int:
        return (n + align - 1) // align * align

    blocks: List[Tuple[str, List[str]]] = []
    for name in order:
        b = block_of(name)
        if not blocks or blocks[-1][0] != b:
            blocks.append((b, []))
        blocks[-1][1].append(name)
    limit = max(int(bucket_mb * (1 << 20)) // 4, 1)
    buckets: List[Bucket] = []
    cur = Bucket(0)
    covered = 0
    for bi, (_, names) in enumerate(blocks):
        last_block = bi == len(blocks) - 1
        if last_block:
            break
        for n in names:
            off, cnt = offsets[n]
            cur.ranges.append((off, padded(cnt)))
            cur.params.append(n)
        cur.trigger = names[-1]
        next_is_last = bi + 1 == len(blocks) - 1
        if cur.elems >= limit or next_is_last or len(buckets) >= max_buckets - 2:
            cur.ranges = _merge(cur.ranges)
            covered += cur.elems
            buckets.append(cur)
            cur = Bucket(len(buckets))
    if cur.ranges:                       # cannot happen (the block before the last always closes), kept for safety
        cur.ranges = _merge(cur.ranges)
        covered += cur.elems
        buckets.append(cur)
        cur = Bucket(len(buckets))
    # final bucket: everything not covered yet
    taken = sorted(r for b in buckets for r in b.ranges)
    rest: List[Tuple[int, int]] = []
    pos = 0
    for off, n in taken:
        if off > pos:
            rest.append((pos, off - pos))
        pos = max(pos, off + n)
    if pos < total_elems:
        rest.append((pos, total_elems - pos))
    cur.ranges = _merge(rest)
    cur.params = [n for n in offsets if all(n not in b.params for b in buckets)]
    cur.final = True
    cur.trigger = None
    buckets.append(cur)
    for b in buckets:
        if len(b.ranges) > MAX_RANGES:
            raise ValueError(f"bucket {b.index} has {len(b.ranges)} ranges (max {MAX_RANGES})")
        for off, n in b.ranges:
            if off % 4 or n % 4:
                raise ValueError("bucket ranges must be multiples of 4 floats")
    assert sum(b.elems for b in buckets) == total_elems, "buckets must tile the flat buffer"
    return buckets


def backward_order(model) -> List[str]:
    """Bulk parameter names of ``models.unet.UNet`` in the order ``NativeEngine.backward`` completes them."""
    bulk = model.flat["bulk_names"]
    specs = model.conv_specs()
    convT = model.up_sample_mode != "bilinear"
    out: List[str] = []

    def add(n: str) -> None:
        if n in bulk:
            out.append(n)

    for i in (1, 2, 3, 4, 5):
        s0, s1 = specs[f"up_conv{i}"]
        add(s1.name + ".weight")
        add(s0.name + ".weight")
        if convT:
            add(f"up_conv{i}.up_sample.weight")
    s0, s1 = specs["double_conv"]
    add(s1.name + ".weight")
    add(s0.name + ".weight")
    for i in (5, 4, 3, 2, 1):
        s0, s1 = specs[f"down_conv{i}"]
        add(s1.name + ".weight")
        add(s0.name + ".weight")
    return out
