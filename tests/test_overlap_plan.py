"""Bucket planning for the gradient exchange that runs under backward (parallel/overlap.py)."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200.models.unet import UNet
from distributed_deep_learning_on_personal_computers_b200.parallel.overlap import (MAX_RANGES, backward_order,
                                                                                     plan_buckets)


@pytest.mark.parametrize("base,mode,mb", [(8, "conv_transpose", 0.01), (16, "bilinear", 0.05), (32, "conv_transpose", 8.0),
                                          (64, "conv_transpose", 8.0), (64, "conv_transpose", 25.0)])
def test_buckets_tile_the_flat_buffer_in_backward_order(base, mode, mb):
    m = UNet(6, mode, base, 3)
    m.flatten_parameters()
    order = backward_order(m)
    total = m.flat["grad"].numel()
    buckets = plan_buckets(m.flat["offsets"], order, total, mb)
    # every element exactly once
    cover = torch.zeros(total, dtype=torch.int32)
    for b in buckets:
        assert 1 <= len(b.ranges) <= MAX_RANGES
        for off, n in b.ranges:
            assert off % 4 == 0 and n % 4 == 0
            cover[off:off + n] += 1
    assert int(cover.min()) == 1 and int(cover.max()) == 1
    # only the last bucket is final; every other one names the wgrad launch that completes it
    assert [b.final for b in buckets] == [False] * (len(buckets) - 1) + [True]
    pos = {n: i for i, n in enumerate(order)}
    last_seen = -1
    for b in buckets[:-1]:
        assert b.trigger in pos
        # the trigger is the LAST parameter of the bucket in backward order, and buckets follow that order
        assert pos[b.trigger] == max(pos[n] for n in b.params)
        assert min(pos[n] for n in b.params) > last_seen
        last_seen = pos[b.trigger]
    # what is only complete when backward ends (BN affine, biases, stem, head) sits in the final bucket
    small = [n for n in m.flat["offsets"] if n not in m.flat["bulk_names"]]
    assert set(small) <= set(buckets[-1].params)
    # the exposed tail stays small: the final bucket holds at most the last block plus the small tensors
    last_block = order[-1].split(".", 1)[0]
    for n in buckets[-1].params:
        assert n in small or n.startswith(last_block)
