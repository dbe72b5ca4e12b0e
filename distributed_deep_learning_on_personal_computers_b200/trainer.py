"""The training loop: micro-batches, local accumulation, gradient exchange, replicated step.

This is the B200-first re-statement of the reference's server / worker main programs
(``R:690-790`` and ``R:792-895``; ``R`` = the reference script).  What is kept:

* every rank holds a full replica and the same optimizer (``R:702-704``, ``R:810``);
* rank 0 is the coordinator: it provides the initial state (``R:708``), owns the run log
  (``R:715-716``, ``R:781-782``) and the sample images (``R:785-790``);
* micro-batches of ``batch_size`` tiles, gradients accumulated locally, exchanged every
  ``sync_every`` micro-batches (``R:685``, ``R:756-763``), then ``optimizer.step()`` on every rank
  with bit-identical gradients (``R:437``, ``R:552``);
* per-epoch mean loss / pixel accuracy and timings (``R:765-782``).

What is different (SURVEY 0.3 decides each quirk): ranks see disjoint shards (Q2); the reduction
is an exact mean (Q1) unless ``reduce="sum"``; accumulated micro-batches are averaged (Q8) unless
``accumulate="sum"``; the tail of an epoch is flushed instead of leaking into the next (Q7);
evaluation runs in ``eval()`` mode on the hold-out split the reference cuts off and never uses
(Q9); data is loaded once (Q12); per-step time is device time (Q13).
"""
from __future__ import annotations

import os
import time
from typing import Dict, Optional

import torch

from .config import TrainConfig
from .models.unet import UNet, count_parameters
from .models.engine import NativeEngine
from .ops import functional as Fn
from .optim import build_optimizer
from .parallel import dist as pdist
from .parallel.allreduce import GradientSynchronizer
from .utils import checkpoint as ckpt
from .utils.data import PinnedBatchLoader, RankSampler, TileSet, build_datasets
from .utils.metrics import DeviceMeter, RunLog, StepReadback, dump_samples
from .utils.timing import PhaseTimer, nvtx_range


class Trainer:
    def __init__(self, cfg: TrainConfig, ctx: Optional[pdist.DistContext] = None,
                 datasets=None):
        self.cfg = cfg
        self.ctx = ctx or pdist.init(cfg)
        cfg.rank, cfg.world_size = self.ctx.rank, self.ctx.world_size
        dev = self.ctx.device
        self.device = dev
        torch.manual_seed(cfg.seed)                       # same seed: ranks agree before broadcast
        self.model = UNet(cfg.out_classes, cfg.up_sample_mode, cfg.base_channels, cfg.in_channels)
        self.model.to(dev)
        self.model.flatten_parameters(dev)
        pdist.broadcast_model_state(self.model, self.ctx)   # rank 0 -> all (R:560-565)

        comm = cfg.comm
        self.star = self.ctx.star if comm == "star" else None    # opened by the rendezvous (parallel/dist.py)
        self.sync = GradientSynchronizer(
            self.model.flat["grad"], self.ctx, codec=cfg.codec, reduce=cfg.reduce,
            accumulate=cfg.accumulate, backend=("auto" if comm == "auto" else comm),
            requantize_output=cfg.requantize_output, star=self.star,
            fused_bulk_elems=self.model.flat.get("bulk_elems", 0), timeout_s=cfg.timeout_s)
        if self.sync.flat is not self.model.flat["grad"]:
            self._rebind_grad(self.sync.flat)

        self.engine_kind = cfg.engine if cfg.engine != "auto" else "native"
        if self.sync.fused and self.ctx.is_coordinator:
            import warnings
            warnings.warn("comm=fused adds weight gradients through the NVSwitch multicast alias: replicas agree to "
                          "rounding (arrival order differs per replica), not bit for bit, and it measured slower than the "
                          "default exchange under backward (DESIGN.md section 4)")
        if self.sync.fused and (self.engine_kind != "native" or os.environ.get("B200SEG_CONV", "") == "cudnn"):
            raise ValueError("comm=fused needs the native engine with the tcgen05 wgrad kernels "
                             "(they are what performs the reduction)")
        self.engine: Optional[NativeEngine] = None
        self._engine_shape = None
        # Default on a multi-GPU node: the exchange runs UNDER backward, bucket by bucket, triggered by the
        # wgrad kernels (parallel/overlap.py); no whole-buffer all-reduce follows the backward pass.
        self.overlap = None
        nat = self.sync.native
        if (cfg.overlap and nat is not None and not self.sync.fused and self.sync.codec.name == "fp32"
                and self.engine_kind == "native" and not cfg.requantize_output):
            from .parallel.overlap import backward_order, plan_buckets
            buckets = plan_buckets(self.model.flat["offsets"], backward_order(self.model),
                                   self.model.flat["grad"].numel(), cfg.bucket_mb)
            nat.setup_overlap(buckets, timeout_s=cfg.timeout_s)
            self.overlap = nat
        self._reduced = False         # the pending gradients were already all-reduced under backward
        self._readback: Optional[StepReadback] = None   # pinned ring of train_step_read
        self.comm_stats: Dict[str, float] = {}
        self.optimizer = build_optimizer(self.model, cfg)
        self.optimizer.rebind_grad(self.model.flat["grad"])

        self.train_set, self.hold_set = datasets if datasets is not None else build_datasets(cfg)
        self.sampler = RankSampler(len(self.train_set), self.ctx.rank, self.ctx.world_size,
                                   cfg.shuffle, cfg.shard_data, cfg.seed)
        self.loader = PinnedBatchLoader(self.train_set, cfg.batch_size, dev)
        self.log = RunLog(cfg, enabled=self.ctx.is_coordinator)
        self.timer = PhaseTimer(dev)
        self.meter = DeviceMeter(dev)
        self.epoch = 0
        self.global_step = 0          # micro-batches seen
        self.pending = 0              # micro-batches accumulated since the last exchange
        self._graphs = {}             # (reduce under backward?) -> (CUDAGraph, launches)
        self.resumed_from: Optional[str] = None
        if cfg.resume:
            # "auto": the newest of <out_dir>/checkpoint.pt and <out_dir>/abort.pt if there is one, else a fresh start --
            # what a supervised job (torchrun --max-restarts N ... --resume auto) needs after PeerLost took it down
            path = ckpt.latest(cfg.out_dir) if cfg.resume == "auto" else cfg.resume
            if path:
                self.resume(path)
                self.resumed_from = path

    # -- plumbing ----------------------------------------------------------------------------------
    def _rebind_grad(self, new_flat: torch.Tensor) -> None:
        """Point every ``.grad`` view (and the engine / optimizer) at ``new_flat``."""
        m = self.model
        convT = {n for n, mod in m.named_modules() if isinstance(mod, torch.nn.ConvTranspose2d)}
        new_flat.zero_()
        for name, p in m.named_parameters():
            off, _ = m.flat["offsets"][name]
            is_t = name.rsplit(".", 1)[0] in convT and p.dim() == 4
            p.grad = UNet._internal_view(new_flat, off, p, is_t)
        m.flat["grad"] = new_flat

    def _ensure_engine(self, n: int, h: int, w: int) -> NativeEngine:
        if self.engine is None or self._engine_shape != (n, h, w):
            self.engine = NativeEngine(self.model, n, h, w)
            self._engine_shape = (n, h, w)
            self.optimizer.attach_lowp(self.engine.flat_lp)
            self._graphs = {}
        return self.engine

    # -- one micro-batch ---------------------------------------------------------------------------
    def micro_step(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, sync_now: bool = False) -> torch.Tensor:
        """Forward + loss + backward; gradients accumulate. Returns device ``[loss_sum, correct]``.

        ``sync_now``: this micro-batch ends an accumulation window.  With the overlapped exchange (the
        default on a multi-GPU node) its backward then also all-reduces the gradient buckets across ranks,
        and the following ``sync_and_step`` only steps the optimizer.
        """
        n, h, w, _ = images_u8.shape
        if self.engine_kind == "torch":
            return self._micro_step_autograd(images_u8, labels_u8)
        eng = self._ensure_engine(n, h, w)
        eng.load_batch(images_u8, labels_u8)
        ov = self.overlap if sync_now else None
        use_graph = self.cfg.cuda_graph and self.device.type == "cuda"
        if not use_graph:
            out = eng.train_microbatch(overlap=ov)
            self._reduced = ov is not None
            return out
        key = ov is not None
        if key not in self._graphs:
            self._graphs[key] = self._capture(eng, ov)
        graph, launches = self._graphs[key]
        with nvtx_range("micro_batch_graph"):
            graph.replay()
        Fn._tick(launches)
        self._reduced = ov is not None
        return eng.loss_acc

    def _capture(self, eng: NativeEngine, ov):
        """Warm up on a side stream (allocator, lazy module loads), then capture one micro-batch.

        The warm-up never touches the fabric: it runs without the overlapped exchange and, in fused mode,
        with the multicast alias unregistered -- a warm-up that added into the peers' buffers would race
        with their own restore of the saved gradients and leave the replicas diverged for good."""
        flat_g = self.model.flat["grad"]
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        saved = flat_g.clone()
        bufs = [b.clone() for b in self.model.buffers()]
        fused = self.sync.fused and self.sync.native is not None
        if fused:
            self.sync.native.suspend_fused()
        with torch.cuda.stream(s):
            eng.train_microbatch()
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        if fused:
            self.sync.native.resume_fused()
        flat_g.copy_(saved)                                  # undo the warm-up's side effects
        for b, keep in zip(self.model.buffers(), bufs):
            b.copy_(keep)
        before = Fn.launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            eng.train_microbatch(overlap=ov)
        launches = Fn.launch_count() - before
        flat_g.copy_(saved)
        for b, keep in zip(self.model.buffers(), bufs):
            b.copy_(keep)
        Fn._tick(-launches)
        return graph, launches

    def train_step(self, images_u8: torch.Tensor, labels_u8: torch.Tensor) -> torch.Tensor:
        """One full data-parallel step (sync every micro-batch): forward, loss, backward with the
        gradient exchange running under it, optimizer.  Returns device ``[loss_sum, correct]``."""
        out = self.micro_step(images_u8, labels_u8, sync_now=True)
        self.pending += 1
        self.global_step += 1
        self.sync_and_step()
        return out

    def train_step_read(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, lag: int = 1):
        """``train_step`` plus a host read of its ``(loss_sum, correct)``, pipelined: this step's 8-byte copy
        is queued behind its kernels, and the call returns the host values of the step ``lag`` calls ago
        (``None`` for the first ``lag`` calls), so the host is never waiting on the step it has just
        launched (the reference's ``.cpu()`` per micro-batch, R:775-776, drains the GPU every time).
        ``flush_reads()`` returns what is still in flight."""
        if self._readback is None:
            self._readback = StepReadback(self.device, depth=max(lag + 1, 2))
        self._readback.push(self.train_step(images_u8, labels_u8))
        return self._readback.pop() if self._readback.in_flight() > lag else None

    def flush_reads(self):
        return [] if self._readback is None else self._readback.drain()

    def _micro_step_autograd(self, images_u8, labels_u8) -> torch.Tensor:
        """Stock-PyTorch path (eager autograd, as the reference's R:754-756), for comparison."""
        x = (images_u8.float() / 255.0).permute(0, 3, 1, 2)
        y = labels_u8.long()
        amp = self.device.type == "cuda" and self.cfg.dtype in ("auto", "bf16")
        if amp:
            x = x.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
            logits = self.model(x)
        loss = torch.nn.functional.cross_entropy(logits.float(), y)
        loss.backward()
        with torch.no_grad():
            px = y.numel()
            correct = (logits.argmax(1) == y).sum()
            return torch.stack([loss.detach() * px, correct.float()])

    # -- exchange + step --------------------------------------------------------------------------
    def sync_and_step(self) -> None:
        k = max(self.pending, 1)
        with self.timer.phase("sync"), nvtx_range("gradient_exchange"):
            if self._reduced:                       # already all-reduced under the backward pass
                scale = self.sync._post_scale(k)
                self._reduced = False
            else:
                scale = self.sync.sync(k)
        with self.timer.phase("optimizer"), nvtx_range("optimizer"):
            self.optimizer.step(grad_scale=scale, zero_grad=True)
            if self.engine is not None and self.optimizer.p_lowp is None:
                self.engine.refresh_lowp()
            self.sync.after_step()
        self.pending = 0

    @torch.no_grad()
    def calibrate_comm(self, images_u8: torch.Tensor, labels_u8: torch.Tensor, reps: int = 3) -> Dict[str, float]:
        """Measure what the gradient exchange costs and how much of it backward hides (metrics.jsonl fields
        ``comm_ms``, ``comm_exposed_ms``, ``overlap_pct``; the reference only prints wall-clock phase times,
        R:265-440).  ``comm_ms``: the whole-buffer all-reduce on its own; ``comm_exposed_ms``: micro-batch with
        the bucketed exchange minus micro-batch without.  Device time, max over ranks.  Leaves parameters,
        gradients, BN buffers and the optimizer untouched.  Collective: every rank must call it."""
        nat = getattr(self.sync, "native", None)
        if nat is None or self.device.type != "cuda":
            return {}
        flat_g = self.model.flat["grad"]
        saved = flat_g.clone()
        bufs = [b.clone() for b in self.model.buffers()]

        def timed(fn) -> float:
            fn()                                           # warm-up (graph capture, lazy loads)
            torch.cuda.synchronize(self.device)
            self.ctx.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize(self.device)
            return pdist.all_reduce_max_time(a.elapsed_time(b) / reps, self.ctx)

        out = {"comm_ms": timed(lambda: nat.all_reduce_())}
        if self.overlap is not None and self.engine_kind == "native":
            def with_ov():
                self.micro_step(images_u8, labels_u8, sync_now=True)
                self._reduced = False
            t_ov = timed(with_ov)
            t_plain = timed(lambda: self.micro_step(images_u8, labels_u8, sync_now=False))
            out["comm_exposed_ms"] = max(t_ov - t_plain, 0.0)
            out["overlap_pct"] = 100.0 * max(0.0, 1.0 - out["comm_exposed_ms"] / max(out["comm_ms"], 1e-9))
        else:
            out["comm_exposed_ms"] = out["comm_ms"]
            out["overlap_pct"] = 0.0
        flat_g.copy_(saved)
        for b, keep in zip(self.model.buffers(), bufs):
            b.copy_(keep)
        self.comm_stats = out
        return out

    def check_peers(self) -> None:
        """Raise ``PeerLost`` if a bounded device-side wait of the gradient exchange gave up (a peer stalled
        or died; the reference would hang forever in ``recv``, R:98).  Rank 0 first writes
        ``<out_dir>/abort.pt`` so the run can be resumed (``--resume``).  Synchronizes the device: called at
        epoch boundaries and every ``check_peers_every`` steps, not per step."""
        nat = getattr(self.sync, "native", None)
        if nat is None:
            return
        from .parallel.star import PeerLost
        try:
            nat.check_abort()
        except PeerLost:
            if self.ctx.is_coordinator:
                try:
                    self.save(os.path.join(self.cfg.out_dir, "abort.pt"))
                except Exception:                      # the abort must surface even if the disk is full
                    pass
            raise

    # -- epochs --------------------------------------------------------------------------------------
    def train_epoch(self) -> Dict[str, float]:
        cfg = self.cfg
        t_epoch = time.perf_counter()
        self.meter.reset()
        self.timer.reset()
        order = self.sampler.indices(self.epoch)
        it = 0
        for images, labels in self.loader.batches(order):
            with self.timer.phase("micro_batch"):
                out = self.micro_step(images, labels, sync_now=((it + 1) % cfg.sync_every == 0))
            self.meter.add(out, labels.numel())
            it += 1
            self.pending += 1
            self.global_step += 1
            if it % cfg.sync_every == 0:                           # R:759
                self.sync_and_step()
                if cfg.check_peers_every and (self.global_step // cfg.sync_every) % cfg.check_peers_every == 0:
                    self.check_peers()
            if cfg.max_steps and self.global_step >= cfg.max_steps:
                break
        if self.pending and cfg.flush_at_epoch_end:                 # Q7
            self.sync_and_step()
        self.check_peers()
        stats = self.meter.read()
        tm = self.timer.summary()
        stats.update(epoch=self.epoch, epoch_s=time.perf_counter() - t_epoch,
                     step_ms=tm.get("micro_batch", 0.0), sync_ms=tm.get("sync", 0.0),
                     optimizer_ms=tm.get("optimizer", 0.0), micro_batches=it)
        if self.device.type == "cuda":
            stats["tiles_per_s"] = cfg.batch_size * 1e3 / max(stats["step_ms"], 1e-9)
        return stats

    @torch.no_grad()
    def evaluate(self, tiles: Optional[TileSet] = None, limit: Optional[int] = None) -> Dict[str, float]:
        """Loss / pixel accuracy in eval mode (running BN statistics) on a tile set."""
        tiles = tiles if tiles is not None else self.hold_set
        if tiles is None or len(tiles) == 0:
            return {}
        n = len(tiles) if limit is None else min(limit, len(tiles))
        b = self.cfg.batch_size
        meter = DeviceMeter(self.device)
        was_training = self.model.training
        self.model.eval()
        for i in range(0, n - b + 1, b):
            img = tiles.images[i:i + b].to(self.device)
            lab = tiles.labels[i:i + b].to(self.device)
            if self.engine_kind == "torch":
                x = (img.float() / 255.0).permute(0, 3, 1, 2)
                logits = self.model(x)
                loss = torch.nn.functional.cross_entropy(logits.float(), lab.long(), reduction="sum")
                out = torch.stack([loss, (logits.argmax(1) == lab.long()).sum().float()])
            else:
                eng = self._ensure_engine(*img.shape[:3])
                eng.load_batch(img, lab)
                out = eng.evaluate()
            meter.add(out, lab.numel())
        self.model.train(was_training)
        r = meter.read()
        return {"eval_loss": r["mean_loss"], "eval_accuracy": r["mean_accuracy"]}

    @torch.no_grad()
    def predict(self, images_u8: torch.Tensor) -> torch.Tensor:
        """Arg-max class map (uint8, N x H x W) for uint8 NHWC tiles, eval mode."""
        img = images_u8.to(self.device)
        if self.engine_kind == "torch":
            self.model.eval()
            out = self.model((img.float() / 255.0).permute(0, 3, 1, 2)).argmax(1).to(torch.uint8)
            self.model.train()
            return out
        preds = []
        b = self.cfg.batch_size
        for i in range(0, img.shape[0], b):
            chunk = img[i:i + b]
            pad = b - chunk.shape[0]
            if pad:
                chunk = torch.cat([chunk, chunk[-1:].expand(pad, -1, -1, -1)])
            eng = self._ensure_engine(*chunk.shape[:3])
            eng.load_batch(chunk, torch.zeros(chunk.shape[:3], dtype=torch.uint8, device=self.device))
            eng.evaluate()
            preds.append(eng.pred[:b - pad].clone())
        return torch.cat(preds)

    def dump_samples(self) -> None:
        k = min(self.cfg.sample_images, len(self.train_set))
        if not self.ctx.is_coordinator or k == 0:
            return
        imgs = self.train_set.images[:k]
        preds = self.predict(imgs)
        dump_samples(os.path.join(self.cfg.out_dir, "samples"), imgs, self.train_set.labels[:k], preds)

    def fit(self) -> Dict[str, float]:
        cfg = self.cfg
        self.log.header(count_parameters(self.model), resumed=self.resumed_from is not None)
        last: Dict[str, float] = {}
        if self.sync.native is not None and len(self.train_set) >= cfg.batch_size:
            b = cfg.batch_size
            self.calibrate_comm(self.train_set.images[:b].to(self.device), self.train_set.labels[:b].to(self.device))
        while self.epoch < cfg.epochs:
            last = self.train_epoch()
            if cfg.eval_holdout and self.hold_set is not None:
                last.update(self.evaluate())
            self.log.epoch(self.epoch, last["mean_loss"], last["mean_accuracy"], last["epoch_s"],
                           last["step_ms"], last["sync_ms"],
                           dict({k: v for k, v in last.items() if k.startswith(("eval_", "tiles_", "optimizer_"))},
                                **self.comm_stats))
            if cfg.sample_images:
                self.dump_samples()
            self.epoch += 1
            if self.ctx.is_coordinator and cfg.save_every_epochs and \
                    self.epoch % cfg.save_every_epochs == 0:
                self.save(os.path.join(cfg.out_dir, "checkpoint.pt"))
            self.ctx.barrier()
            if cfg.max_steps and self.global_step >= cfg.max_steps:
                break
        return last

    # -- checkpoint --------------------------------------------------------------------------------
    def save(self, path: str) -> str:
        return ckpt.save(path, self.model, self.optimizer, self.epoch, self.global_step,
                         self.cfg.to_dict())

    def resume(self, path: str) -> None:
        state = ckpt.load(path, self.model, self.optimizer)
        self.epoch, self.global_step = state["epoch"], state["step"]
        if self.engine is not None:
            self.engine.refresh_lowp()

    def close(self) -> None:
        if getattr(self.sync, "native", None) is not None:
            self.sync.native.close()
        self.ctx.shutdown()
