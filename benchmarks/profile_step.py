#!/usr/bin/env python
"""Kernel-level breakdown of one training step (torch.profiler / CUPTI; times are for SHARES only).

    python benchmarks/profile_step.py --impl own|reference [--conv tcgen05|cudnn] [--out FILE]
"""
import argparse
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="own")
    ap.add_argument("--conv", default="tcgen05")
    ap.add_argument("--base-channels", type=int, default=64)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--tile", type=int, default=512)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    ap.add_argument("--comm", default="auto", help="under torchrun: native | fused | torch")
    ap.add_argument("--graph", action="store_true", help="profile CUDA-graph replays (steady state, as bench.py runs)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    os.environ["B200SEG_CONV"] = a.conv
    from distributed_deep_learning_on_personal_computers_b200.utils.data import synthetic_tiles
    tiles = synthetic_tiles(a.batch, a.tile, 6, 0)
    if a.impl == "own":
        from distributed_deep_learning_on_personal_computers_b200 import TrainConfig
        from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer
        cfg = TrainConfig(base_channels=a.base_channels, tile=a.tile, batch_size=a.batch, sync_every=1, codec="fp32",
                          world_size=world, rank=rank, comm=a.comm, train_tiles=a.batch, holdout_tiles=0,
                          sample_images=0, cuda_graph=a.graph,
                          out_dir=os.path.join(ROOT, "gpurun_out", "prof_run"), device="cuda")
        if world > 1:
            cfg = cfg.apply_env()
        tr = Trainer(cfg, datasets=(tiles, None))
        img, lab = tiles.images.cuda(), tiles.labels.cuda()

        def step():
            tr.micro_step(img, lab)
            tr.pending = 1
            tr.sync_and_step()
    else:
        from baseline.ref_loader import reference_unet
        model = reference_unet(a.base_channels, 6).cuda().to(memory_format=torch.channels_last)
        opt = torch.optim.Adam(model.parameters())
        crit = torch.nn.CrossEntropyLoss()
        torch.backends.cudnn.benchmark = True
        x = (tiles.images.float() / 255).permute(0, 3, 1, 2).cuda().contiguous(memory_format=torch.channels_last)
        y = tiles.labels.long().cuda()

        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = crit(model(x).float(), y)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        step()
    e1.record()
    torch.cuda.synchronize()
    step_ms = e0.elapsed_time(e1) / 10
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(a.steps):
            step()
        torch.cuda.synchronize()
    rows = []
    for ev in prof.key_averages():
        t = getattr(ev, "device_time_total", None) or getattr(ev, "cuda_time_total", 0)
        if t:
            rows.append((t / a.steps / 1e3, ev.count / a.steps, ev.key))
    rows.sort(reverse=True)
    total = sum(r[0] for r in rows)
    lines = [f"total kernel ms/step {total:.3f}  impl={a.impl} conv={a.conv}  graph={getattr(a, 'graph', False)}  "
             f"step (CUDA events, no profiler) {step_ms:.3f} ms -> not in kernels {step_ms - total:.3f} ms"]
    for ms, cnt, name in rows[:40]:
        lines.append(f"{ms:9.3f} ms {100 * ms / total:5.1f}%  x{cnt:6.1f}  {name[:110]}")
    text = "\n".join(lines)
    if rank == 0:
        print(text)
        if a.out:
            open(a.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
