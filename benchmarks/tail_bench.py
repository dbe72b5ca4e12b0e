#!/usr/bin/env python
"""Time the network's tail at the headline shape: three kernels (BN+ReLU forward, head + loss + backward,
BN-backward reduce) against the one-pass kernel of csrc/elem/head_tail.cu.  CUDA events, L2 flushed by size
(each tensor is 537 MB).  ``python benchmarks/tail_bench.py [--n 16 --tile 512 --c 64]``"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn  # noqa: E402


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16)
    ap.add_argument("--tile", type=int, default=512)
    ap.add_argument("--c", type=int, default=64)
    ap.add_argument("--classes", type=int, default=6)
    args = ap.parse_args()
    dev = "cuda"
    N, H, C, K = args.n, args.tile, args.c, args.classes
    torch.manual_seed(0)
    y = (torch.randn(N, H, H, C, device=dev) * 1.5).to(torch.bfloat16)
    ss = torch.stack([torch.rand(C) + 0.5, torch.randn(C) * 0.3, torch.randn(C) * 0.1, torch.rand(C) + 0.5]).to(dev)
    w, b = torch.randn(K, C, device=dev) * 0.2, torch.randn(K, device=dev) * 0.1
    lab = torch.randint(0, K, (N, H, H), dtype=torch.uint8, device=dev)
    gamma = torch.rand(C, device=dev) + 0.5
    a, g, dy = torch.empty_like(y), torch.empty_like(y), torch.empty_like(y)
    dw, db = torch.zeros(K, C, device=dev), torch.zeros(K, device=dev)
    dg, dbt = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    pred = torch.empty(N, H, H, dtype=torch.uint8, device=dev)
    acc = torch.zeros(2, device=dev)
    rows = Fn.head_tail_rows(y)
    part = torch.zeros(rows, 2, C, device=dev)
    scale = 1.0 / (N * H * H)
    out = {"shape": [N, H, H, C], "classes": K}
    out["bn_relu_fwd_ms"] = timeit(lambda: Fn.bn_relu_pool_fwd(y, ss, a, None))
    out["head_ce_ms"] = timeit(lambda: Fn.head_ce_fwd_bwd(a, w, b, lab, scale, g, dw, db, pred, acc))
    out["bn_bwd_3kernel_ms"] = timeit(lambda: Fn.bn_relu_pool_bwd(g, None, y, ss, gamma, dg, dbt, None, dy))
    loss_old = acc.clone()
    out["head_tail_ms"] = timeit(lambda: Fn.head_tail_fwd_bwd(y, ss, w, b, lab, scale, g, dw, db, pred, acc, part))
    out["bn_bwd_from_partial_ms"] = timeit(lambda: Fn.bn_relu_bwd_from_partial(part, rows, g, y, ss, gamma, dg, dbt, dy))
    out["old_chain_ms"] = out["bn_relu_fwd_ms"] + out["head_ce_ms"] + out["bn_bwd_3kernel_ms"]
    out["new_chain_ms"] = out["head_tail_ms"] + out["bn_bwd_from_partial_ms"]
    gb = y.numel() * 2 * 2 / 1e9                          # one read of y, one write of g
    out["head_tail_TBps"] = gb / out["head_tail_ms"]
    out["loss_old_new"] = [loss_old[0].item(), acc[0].item()]
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in out.items()}))


if __name__ == "__main__":
    main()
