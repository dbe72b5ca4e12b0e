"""Run log, JSONL metrics, sample-image dumps and phase timers.

Reference behaviour (``R`` = the reference script):

* run log ``otus_{model_bytes}.txt`` (``R:715-716``): a header with the per-PC batch, the global
  batch ``batch_size * (N_conn + 1)``, the sync period, the width divisor and the PC count, then
  one line per epoch with epoch, mean loss, mean pixel accuracy and timings (``R:781-782``);
* per-epoch metrics: running mean of the loss and of the pixel accuracy over the micro-batches
  (``R:775-777``), read back from the device every micro-batch (a forced sync);
* sample dump: for tiles 0..4 ``Model i.png`` (arg-max * 5 through matplotlib's default colour
  map), ``Label i.png`` (label * 5) and ``Image i.png`` (``R:785-790``), overwritten every epoch;
* phase timings printed from ``time.time()`` deltas around each step of the sync (C27).

Here the loss / correct-pixel counters stay on the device and are read once per logging
interval; PNGs are written with PIL only (matplotlib is not needed; a small viridis-like table
stands in for its default colour map); timings come from CUDA events when on a GPU; everything
is also appended to ``metrics.jsonl`` for machines to read.  Only rank 0 writes.
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional

import numpy as np
import torch

# Seven anchors of matplotlib's default (viridis) map, linearly interpolated to 256 entries.
_ANCHORS = np.array([[68, 1, 84], [70, 50, 127], [54, 92, 141], [39, 127, 142],
                     [31, 161, 135], [74, 194, 109], [160, 218, 57], [253, 231, 37]], np.float64)


def _viridis_lut() -> np.ndarray:
    xs = np.linspace(0, len(_ANCHORS) - 1, 256)
    lo = np.floor(xs).astype(int).clip(0, len(_ANCHORS) - 2)
    f = (xs - lo)[:, None]
    return (_ANCHORS[lo] * (1 - f) + _ANCHORS[lo + 1] * f).round().astype(np.uint8)


_LUT = _viridis_lut()


def colorize(classmap: np.ndarray) -> np.ndarray:
    """What ``matplotlib.pyplot.imsave(uint8(map * 5))`` does (R:788): min-max normalise, then map."""
    v = classmap.astype(np.float64) * 5.0
    lo, hi = float(v.min()), float(v.max())
    norm = np.zeros_like(v) if hi <= lo else (v - lo) / (hi - lo)
    return _LUT[(norm * 255).round().astype(np.uint8)]


def dump_samples(out_dir: str, images_u8: torch.Tensor, labels: torch.Tensor,
                 preds: torch.Tensor) -> List[str]:
    """Write ``Model i.png`` / ``Label i.png`` / ``Image i.png`` triplets (R:785-790)."""
    from PIL import Image
    os.makedirs(out_dir, exist_ok=True)
    written = []
    for i in range(images_u8.shape[0]):
        trip = {f"Model {i}.png": colorize(preds[i].cpu().numpy()),
                f"Label {i}.png": colorize(labels[i].cpu().numpy()),
                f"Image {i}.png": images_u8[i].cpu().numpy()}
        for name, arr in trip.items():
            path = os.path.join(out_dir, name)
            Image.fromarray(arr).save(path)
            written.append(path)
    return written


class RunLog:
    """Reference-style text log plus JSONL (rank 0 only)."""

    def __init__(self, cfg, enabled: bool = True):
        self.enabled = enabled
        self.cfg = cfg
        self.path = cfg.log_file()
        self.jsonl = os.path.join(cfg.out_dir, "metrics.jsonl")
        if enabled:
            os.makedirs(cfg.out_dir, exist_ok=True)

    def header(self, n_params: int, resumed: bool = False) -> None:
        if not self.enabled:
            return
        c = self.cfg
        line = (f"batch_per_rank={c.batch_size} global_batch={c.global_batch} "
                f"sync_every={c.sync_every} base_channels={c.base_channels} "
                f"width_divisor={64 // c.base_channels if c.base_channels <= 64 else 0} "
                f"ranks={c.world_size} codec={c.codec} comm={c.comm} params={n_params}")
        with open(self.path, "a" if resumed else "w") as fh:
            fh.write(("# resumed\n" if resumed else "") + line + "\n")
        self.event({"event": "header", "config": c.to_dict(), "params": n_params})

    def epoch(self, epoch: int, mean_loss: float, mean_acc: float, epoch_s: float,
              step_ms: float, sync_ms: float, extra: Optional[Dict] = None) -> None:
        if not self.enabled:
            return
        with open(self.path, "a") as fh:
            fh.write(f"epoch={epoch} mean_loss={mean_loss:.6f} mean_accuracy={mean_acc:.6f} "
                     f"epoch_s={epoch_s:.3f} step_ms={step_ms:.3f} sync_ms={sync_ms:.3f}\n")
        rec = {"event": "epoch", "epoch": epoch, "mean_loss": mean_loss, "mean_accuracy": mean_acc,
               "epoch_s": epoch_s, "step_ms": step_ms, "sync_ms": sync_ms}
        rec.update(extra or {})
        self.event(rec)

    def event(self, rec: Dict) -> None:
        if not self.enabled:
            return
        rec = dict(rec, time=time.time())
        with open(self.jsonl, "a") as fh:
            fh.write(json.dumps(rec, default=str) + "\n")


class DeviceMeter:
    """Accumulates ``[loss_sum, correct]`` pairs on the device; one read-back per interval."""

    def __init__(self, device: torch.device):
        self.acc = torch.zeros(2, dtype=torch.float64, device=device)
        self.pixels = 0

    def add(self, loss_correct: torch.Tensor, pixels: int) -> None:
        self.acc += loss_correct.to(torch.float64)
        self.pixels += pixels

    def read(self) -> Dict[str, float]:
        vals = self.acc.tolist()              # the only D2H sync
        px = max(self.pixels, 1)
        return {"mean_loss": vals[0] / px, "mean_accuracy": vals[1] / px, "pixels": self.pixels}

    def reset(self) -> None:
        self.acc.zero_()
        self.pixels = 0


class StepReadback:
    """Per-step device-to-host read of ``[loss_sum, correct]`` that never stalls the step it belongs to.

    The reference reads loss and accuracy back with ``.cpu()`` right after every micro-batch (R:775-776):
    the host then cannot queue the next step until this one has drained, and the GPU idles for the host's
    launch time at every step.  Here ``push`` queues the 8-byte copy into a pinned ring slot behind the
    step's kernels and records an event; ``pop`` waits for the OLDEST outstanding slot.  Called as
    ``push(step i); pop() -> values of step i-1`` the host waits while the GPU is already running step i,
    so every step's result still reaches the host, one step late."""

    def __init__(self, device: torch.device, depth: int = 4, width: int = 2):
        self.cuda = device.type == "cuda"
        self.slots = [torch.zeros(width, dtype=torch.float32, pin_memory=self.cuda) for _ in range(depth)]
        self.events = [torch.cuda.Event() for _ in range(depth)] if self.cuda else None
        self.head = self.tail = 0               # tail .. head-1 are in flight
        self.bytes_per_read = width * 4

    def in_flight(self) -> int:
        return self.head - self.tail

    def push(self, values: torch.Tensor) -> None:
        if self.in_flight() == len(self.slots):
            raise RuntimeError("StepReadback ring is full: pop() before pushing more")
        k = self.head % len(self.slots)
        self.slots[k].copy_(values.detach().to(torch.float32), non_blocking=True)
        if self.cuda:
            self.events[k].record()
        self.head += 1

    def pop(self):
        """Host values of the oldest pushed step (waits for it), or None when nothing is in flight."""
        if self.in_flight() == 0:
            return None
        k = self.tail % len(self.slots)
        if self.cuda:
            self.events[k].synchronize()
        self.tail += 1
        return tuple(self.slots[k].tolist())

    def drain(self):
        out = []
        while self.in_flight():
            out.append(self.pop())
        return out
