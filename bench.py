#!/usr/bin/env python
"""Headline benchmark: U-Net train tiles/s (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5
    python bench.py --impl reference ...      # the reference's own UNet under stock DDP/NCCL/cuDNN

Config (BASELINE.json config 2): U-Net with 64 base channels (the reference's ``NN_in_model = 1``;
config 3 calls 128 "2x width"), 512x512x3 tiles, 6 classes, batch 16 per GPU, bf16 compute, Adam,
one gradient exchange + optimizer step per micro-batch (DDP semantics), synthetic Vaihingen-shaped
data, random init.  A "step" = forward + loss + backward + gradient all-reduce + optimizer step.

Two timed regions, both with CUDA events bracketed by barrier + synchronize, max over ranks:
  * ``value``     -- the batch is already resident on the device (kernels + collective only);
  * ``e2e.value`` -- through the public API (``Trainer`` + ``PinnedBatchLoader``): every step copies
                     its uint8 tiles from pinned host memory and reads the loss back to the host.
Working set per step is several GB of activations (>> 126 MB L2), so no explicit L2 flush.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_TILES_PER_S = None      # BASELINE.md publishes no throughput number -> vs_baseline null
# BASELINE.json configs (config 5, the all-reduce sweep, is benchmarks/allreduce_sweep.py)
PRESETS = {
    "headline": {"base_channels": 64, "tile": 512, "batch": 16},
    "b128": {"base_channels": 128, "tile": 512, "batch": 16},
    "tile1024": {"base_channels": 64, "tile": 1024, "batch": 2},
    "shipped": {"base_channels": 32, "tile": 512, "batch": 1},
}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="own", choices=["own", "reference"])
    p.add_argument("--base-channels", type=int, default=64)
    p.add_argument("--tile", type=int, default=512)
    p.add_argument("--batch", type=int, default=16)
    p.add_argument("--codec", default="fp32")
    p.add_argument("--comm", default="auto")
    p.add_argument("--conv", default=None, help="tcgen05 | cudnn (own arm only)")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--readback", choices=["lagged", "sync"], default="lagged",
                   help="e2e arm: host picks up step i's loss while step i+1 runs (lagged) or waits for it at once (sync)")
    p.add_argument("--no-overlap", action="store_true", help="exchange after backward (whole buffer) instead of under it")
    p.add_argument("--preset", default=None, choices=sorted(PRESETS),
                   help="BASELINE.json configs: headline (2, default), b128 (3), tile1024 (4), shipped (reference as shipped)")
    a = p.parse_args()
    if a.preset:
        for k, v in PRESETS[a.preset].items():
            setattr(a, k, v)
    return a


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs (rank 0)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(float(r[1])) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def timed(fn, steps, dist_ctx, torch, finish=None):
    """Device time of ``steps`` calls of fn (then ``finish``, inside the timed region), max over ranks, in ms."""
    import torch.distributed as dist
    if dist_ctx["world"] > 1:
        dist.barrier(device_ids=[dist_ctx["local"]])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        fn(i)
    if finish is not None:
        finish()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    if dist_ctx["world"] > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier(device_ids=[dist_ctx["local"]])
        ms = float(t.item())
    return ms


def _quiet_stdout():
    """Route fd 1 to stderr while the benchmark runs (NCCL and friends print banners on stdout) and
    return a writer for the ONE JSON line the driver expects on the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)

    def emit(line: str) -> None:
        sys.stdout.flush()
        os.write(real, (line + "\n").encode())
    return emit


def main():
    args = parse()
    emit = _quiet_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    if not torch.cuda.is_available():
        if rank == 0:
            emit(json.dumps({"impl": args.impl, "unavailable": "no CUDA device on this machine"}))
        return 0
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dctx = {"world": world, "rank": rank, "local": local}
    if args.impl == "reference":
        out = run_reference(args, dctx, torch)
    else:
        out = run_own(args, dctx, torch)
    if rank == 0:
        emit(json.dumps(out))
    if world > 1:
        import torch.distributed as dist
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()
    return 0


def _result(args, dctx, impl, ms, ms_e2e, clocks, launches, h2d, d2h, extra_cfg):
    world = dctx["world"]
    tiles = args.batch * world * args.steps
    val = tiles / (ms / 1e3)
    out = {"metric": "U-Net train tiles/sec", "value": round(val, 3), "unit": "tiles/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": (val / BASELINE_TILES_PER_S) if BASELINE_TILES_PER_S else None,
           "dtype": "bf16", "data": "synthetic", "impl": impl,
           "config": dict({"model": f"unet_b{args.base_channels}", "global_batch": args.batch * world,
                           "seq_len": args.tile, "tile": f"{args.tile}x{args.tile}x3", "classes": 6,
                           "batch_per_gpu": args.batch, "parallelism": f"dp{world}",
                           "optimizer": "adam", "sync_every": 1,
                           "l2": "working set per step >> 126 MB L2 (no explicit flush)"},
                          **extra_cfg),
           "clocks": clocks, "gpu_launches": launches}
    if ms_e2e is not None:
        out["e2e"] = {"value": round(tiles / (ms_e2e / 1e3), 3), "unit": "tiles/s",
                      "ms_per_step": round(ms_e2e / args.steps, 4),
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}
    return out


# ------------------------------------------------------------------------------------------------
def run_own(args, dctx, torch):
    from distributed_deep_learning_on_personal_computers_b200 import TrainConfig
    from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer
    from distributed_deep_learning_on_personal_computers_b200.parallel import dist as pdist
    from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn
    from distributed_deep_learning_on_personal_computers_b200.utils.data import synthetic_tiles

    if args.conv:
        os.environ["B200SEG_CONV"] = args.conv
    world, rank, local = dctx["world"], dctx["rank"], dctx["local"]
    n_tiles = args.batch * 3
    cfg = TrainConfig(base_channels=args.base_channels, tile=args.tile, batch_size=args.batch,
                      sync_every=1, codec=args.codec, comm=args.comm, world_size=world, rank=rank,
                      local_rank=local, train_tiles=n_tiles, holdout_tiles=0, sample_images=0,
                      cuda_graph=not args.no_graph, out_dir=os.path.join(ROOT, "gpurun_out", "bench_run"),
                      shard_data=False, shuffle=False, device="cuda", overlap=not args.no_overlap,
                      check_peers_every=0)
    ctx = pdist.DistContext(rank, world, local, torch.device("cuda", local),
                            "nccl" if world > 1 else "none", False)
    tiles = synthetic_tiles(n_tiles, args.tile, 6, seed=rank)
    tr = Trainer(cfg, ctx=ctx, datasets=(tiles, None))
    dev_batches = [(tiles.images[i * args.batch:(i + 1) * args.batch].cuda(),
                    tiles.labels[i * args.batch:(i + 1) * args.batch].cuda()) for i in range(3)]

    def step_resident(i):
        img, lab = dev_batches[i % 3]
        tr.train_step(img, lab)       # forward + loss + backward (gradient buckets reduced under it) + Adam

    for i in range(max(args.warmup, 3)):
        step_resident(i)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    Fn.reset_launch_count()
    ms = timed(step_resident, args.steps, dctx, torch)
    launches = Fn.launch_count()
    clk = clocks.stop() if rank == 0 else None

    ms_e2e, h2d, d2h = None, 0, 0
    if not args.no_e2e:
        order = [j % n_tiles for j in range(args.batch * (args.steps + max(args.warmup, 3)))]
        it = tr.loader.batches(order)
        reads = []

        def step_e2e(i):
            img, lab = next(it)                      # H2D from pinned memory on the copy stream
            # every step's (loss_sum, correct) is copied to pinned host memory behind its kernels; the host
            # picks it up one call later, while the next step runs (args.readback sync: wait for it at once)
            got = tr.train_step_read(img, lab, lag=0 if args.readback == "sync" else 1)
            if got is not None:
                reads.append(got)

        for i in range(max(args.warmup, 3)):
            step_e2e(i)
        tr.flush_reads()
        reads.clear()
        # the last step's values are read inside the timed region too (finish runs before the closing event)
        ms_e2e = timed(step_e2e, args.steps, dctx, torch, finish=lambda: reads.extend(tr.flush_reads()))
        assert len(reads) == args.steps and all(v[0] == v[0] for v in reads), "a step's result was not read back"
        h2d, d2h = tr.loader.bytes_per_batch, 8
    tr.check_peers()
    extra = {"engine": "native", "conv": os.environ.get("B200SEG_CONV", "tcgen05"),
             "comm": ("fused" if tr.sync.fused else tr.sync.backend), "codec": args.codec, "cuda_graph": not args.no_graph,
             "allreduce": (None if tr.sync.native is None else
                           (("nvls" if tr.overlap.multicast_ptr else "twoshot") + "-bucketed-under-backward"
                            if tr.overlap is not None else tr.sync.native.algorithm)),
             "grad_buckets": (len(tr.overlap.ovl_buckets) if tr.overlap is not None else 0),
             "e2e_readback": ("every step, picked up one step late" if args.readback == "lagged" else "every step, waited for at once"),
             # every B200SEG_* switch in effect (they select kernel variants; recorded so no run is ambiguous)
             "env": {k: v for k, v in sorted(os.environ.items()) if k.startswith("B200SEG_")}}
    return _result(args, dctx, "own", ms, ms_e2e, clk, launches, h2d, d2h, extra)


# ------------------------------------------------------------------------------------------------
def run_reference(args, dctx, torch):
    """The reference's own ``UNet`` class (R:575-656, loaded unmodified from baseline/_ref) under
    stock DistributedDataParallel + NCCL + cuDNN/cuBLAS, bf16 autocast, channels-last, Adam."""
    try:
        from baseline.ref_loader import reference_unet
        model = reference_unet(args.base_channels, 6).cuda()
    except Exception as exc:
        return {"impl": "reference", "unavailable": f"{type(exc).__name__}: {exc}"[:200]}
    from distributed_deep_learning_on_personal_computers_b200.utils.data import synthetic_tiles
    world, rank, local = dctx["world"], dctx["rank"], dctx["local"]
    torch.backends.cudnn.benchmark = True
    model = model.to(memory_format=torch.channels_last)
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP
        model = DDP(model, device_ids=[local], gradient_as_bucket_view=True)
    opt = torch.optim.Adam(model.parameters())               # R:704
    crit = torch.nn.CrossEntropyLoss()                         # R:703
    n_tiles = args.batch * 3
    tiles = synthetic_tiles(n_tiles, args.tile, 6, seed=rank)
    # The reference keeps float32 NCHW images and int64 labels on the host (R:737-741).
    x_host = (tiles.images.float() / 255).permute(0, 3, 1, 2).contiguous().pin_memory()
    y_host = tiles.labels.long().pin_memory()
    B = args.batch
    dev = [(x_host[i * B:(i + 1) * B].cuda().contiguous(memory_format=torch.channels_last),
            y_host[i * B:(i + 1) * B].cuda()) for i in range(3)]

    def train_step(x, y):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
            loss = crit(out.float(), y)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    def step_resident(i):
        x, y = dev[i % 3]
        train_step(x, y)

    for i in range(max(args.warmup, 3)):
        step_resident(i)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(step_resident, args.steps, dctx, torch)
    clk = clocks.stop() if rank == 0 else None

    ms_e2e, h2d, d2h = None, 0, 0
    if not args.no_e2e:
        def step_e2e(i):
            j = i % 3
            x = x_host[j * B:(j + 1) * B].cuda(non_blocking=True).contiguous(memory_format=torch.channels_last)
            y = y_host[j * B:(j + 1) * B].cuda(non_blocking=True)
            loss = train_step(x, y)
            loss.item()                                     # R:776 reads the loss every micro-batch
        for i in range(max(args.warmup, 3)):
            step_e2e(i)
        ms_e2e = timed(step_e2e, args.steps, dctx, torch)
        h2d = B * args.tile * args.tile * (3 * 4 + 8)
        d2h = 4
    extra = {"engine": "torch eager autograd + DDP", "conv": "cudnn", "comm": "nccl" if world > 1 else "none"}
    return _result(args, dctx, "reference", ms, ms_e2e, clk, 0, h2d, d2h, extra)


if __name__ == "__main__":
    sys.exit(main())
