#!/usr/bin/env python
"""Gradient all-reduce bandwidth sweep, 64 KB - 512 MB: our peer-memory kernels vs NCCL (BASELINE config 5).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        benchmarks/allreduce_sweep.py [--max-mb 512] [--out FILE]

Device-timed (CUDA events), 5 warm-up + 20 timed calls per size, max over ranks.  Bus bandwidth =
2 (W-1)/W * bytes / time, the NCCL convention.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_deep_learning_on_personal_computers_b200.parallel import symm, dist as pdist  # noqa: E402
from distributed_deep_learning_on_personal_computers_b200.parallel.codecs import get_codec  # noqa: E402


def timed(fn, iters, world, dev):
    for _ in range(5):
        fn()
    dist.barrier(device_ids=[dev])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / iters], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=512)
    ap.add_argument("--out", default="")
    ap.add_argument("--grid-sweep", action="store_true", help="also time the NVLS kernel at 32..256 CTAs")
    ap.add_argument("--min-mb", type=float, default=0.0625)
    ap.add_argument("--nvls-only", action="store_true")
    ap.add_argument("--blocks", default="32,64,128,256")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = pdist.DistContext(rank, world, local, torch.device("cuda", local), "nccl", False)
    rows = []
    size = int(a.min_mb * (1 << 20))
    while size <= a.max_mb << 20:
        n = size // 4
        x = torch.randn(n, device="cuda")
        row = {"bytes": size, "world": world}
        row["nccl_ms"] = timed(lambda: dist.all_reduce(x), 20, world, local)
        # lossy codecs go through the quantize-on-send push kernel (amax included for the scaled ones)
        variants = [("twoshot", "fp32"), ("twoshot", "bf16"), ("twoshot", "int8s")]
        if size <= symm.ONESHOT_MAX_BYTES:
            variants.append(("oneshot", "fp32"))
        if a.nvls_only:
            variants = []
        probe = symm.SymmetricAllReduce(x, ctx, get_codec("fp32"), algorithm="twoshot")
        if probe.multicast_ptr:
            variants.append(("nvls", "fp32"))
        del probe
        for algo, codec in variants:
            try:
                ar = symm.SymmetricAllReduce(x, ctx, get_codec(codec), algorithm=algo)
                row[f"{algo}_{codec}_ms"] = timed(lambda: ar.all_reduce_(), 20, world, local)
                if algo == "nvls" and a.grid_sweep and size >= (16 << 20):
                    for blocks in [int(b) for b in a.blocks.split(",")]:      # CTAs of 512 threads
                        ar.blocks = blocks
                        row[f"nvls_fp32_b{blocks}_ms"] = timed(lambda: ar.all_reduce_(), 20, world, local)
                del ar
            except Exception as exc:                      # keep sweeping
                row[f"{algo}_{codec}_err"] = str(exc)[:80]
        bus = 2 * (world - 1) / world * size
        for k in list(row):
            if k.endswith("_ms"):
                row[k.replace("_ms", "_busGBs")] = round(bus / row[k] / 1e6, 1)
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        size *= 4
    if rank == 0 and a.out:
        with open(a.out, "w") as fh:
            for r in rows:
                fh.write(json.dumps(r) + "\n")
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
