#!/usr/bin/env python
"""Tiny driver for ncu: a few training micro-batches of the flagship config, no CUDA graph.

    python benchmarks/ncu_target.py [--steps 2] [--base-channels 64] [--batch 16] [--tile 512]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_deep_learning_on_personal_computers_b200 import TrainConfig  # noqa: E402
from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer  # noqa: E402
from distributed_deep_learning_on_personal_computers_b200.utils.data import synthetic_tiles  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--base-channels", type=int, default=64)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--tile", type=int, default=512)
a = ap.parse_args()
tiles = synthetic_tiles(a.batch, a.tile, 6, 0)
cfg = TrainConfig(base_channels=a.base_channels, tile=a.tile, batch_size=a.batch, sync_every=1, codec="fp32",
                  world_size=1, train_tiles=a.batch, holdout_tiles=0, sample_images=0, cuda_graph=False,
                  out_dir=os.path.join(ROOT, "gpurun_out", "ncu_run"), device="cuda")
tr = Trainer(cfg, datasets=(tiles, None))
img, lab = tiles.images.cuda(), tiles.labels.cuda()
for i in range(a.steps):
    last = i == a.steps - 1
    if last:                                   # ncu --profile-from-start off captures exactly this step
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    tr.micro_step(img, lab)
    tr.pending = 1
    tr.sync_and_step()
    if last:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
torch.cuda.synchronize()
print("ok")
