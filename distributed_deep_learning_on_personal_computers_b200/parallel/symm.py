"""Symmetric-memory all-reduce: the Python side of ``csrc/comm/allreduce.cu``.

The reference's rendezvous opens one TCP connection per worker (``create_server`` /
``create_worker``, ``R:173-206``) and every later exchange goes through the root.  Here the
rendezvous exchanges *memory handles*: each rank allocates its gradient buffer, its wire buffer
and a signal pad from CUDA's peer-shareable allocator (``torch.distributed._symmetric_memory``
does the handle exchange over the existing process group) and maps every peer's copies into its
own address space.  After that the gradient exchange is one kernel launch per step with no host
involvement, no NCCL, and no root.

``SymmetricAllReduce.buffer`` is the flat fp32 gradient buffer the engine accumulates into
(``Trainer`` re-homes ``.grad`` there).  ``all_reduce_()`` leaves the rank-order fp32 sum of all
replicas in it on every rank, bit-identical.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from .codecs import Codec
from .dist import DistContext

_WIRE_KIND = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.int8: 3}
ONESHOT_MAX_BYTES = 512 * 1024


def available(ctx: DistContext) -> bool:
    """Native path needs CUDA, NCCL process group (for the handle exchange), P2P between all ranks
    of the node, and the library."""
    if ctx.device.type != "cuda" or ctx.world_size < 2 or ctx.world_size > 16:
        return False
    from ..ops import loader
    if loader.library() is None:
        return False
    try:
        import torch.distributed._symmetric_memory  # noqa: F401
    except Exception:
        return False
    n = torch.cuda.device_count()
    if n < ctx.world_size:
        return False
    me = ctx.device.index
    return all(torch.cuda.can_device_access_peer(me, d) for d in range(n) if d != me)


class SymmetricAllReduce:
    def __init__(self, like_flat: torch.Tensor, ctx: DistContext, codec: Codec, group=None,
                 algorithm: str = "auto", fused_bulk_elems: int = 0, timeout_s: float = 600.0):
        import torch.distributed._symmetric_memory as symm_mem
        from ..ops import loader
        if loader.library() is None:
            raise RuntimeError(f"native all-reduce needs the b200seg library: {loader.load_error()}")
        self.ctx, self.codec = ctx, codec
        self.group = group or dist.group.WORLD
        dev = like_flat.device
        n = like_flat.numel()
        self.n = (n + 63) // 64 * 64
        self.buffer_full = symm_mem.empty(self.n, dtype=torch.float32, device=dev)
        self.buffer_full.zero_()
        self.buffer = self.buffer_full[:n]
        self.h_grad = symm_mem.rendezvous(self.buffer_full, self.group)
        ops = torch.ops.b200seg
        self.timeout_s = float(timeout_s)
        ops.set_comm_timeout(self.timeout_s)               # bounds every device-side wait for a peer
        self.abort_word = int(ops.allreduce_abort_word())
        sig_words = ops.allreduce_signal_bytes() // 4 + 64
        self.signal = symm_mem.empty(sig_words, dtype=torch.int32, device=dev)
        self.signal.zero_()
        self.h_sig = symm_mem.rendezvous(self.signal, self.group)
        self.mult_off = ops.allreduce_signal_bytes()            # float slot after the flags
        self.mult_slot = self.signal.view(torch.float32)[self.mult_off // 4:self.mult_off // 4 + 1]
        self.mult_slot.fill_(1.0)
        self.wire = None
        self.h_wire = None
        fp8 = getattr(torch, "float8_e4m3fn", None)
        if codec.name != "fp32":
            # + slack: the push kernel's inbox holds world slots of ceil(vectors / world) 16-byte vectors each
            self.wire = symm_mem.empty(self.n * codec.wire_bytes + 4096, dtype=torch.uint8, device=dev)
            self.wire.zero_()
            self.h_wire = symm_mem.rendezvous(self.wire, self.group)
            self.wire_kind = 4 if (fp8 is not None and codec.wire_dtype == fp8) else _WIRE_KIND[codec.wire_dtype]
        else:
            self.wire_kind = 0
        self.grad_ptrs = [int(p) for p in self.h_grad.buffer_ptrs]
        self.sig_ptrs = [int(p) for p in self.h_sig.buffer_ptrs]
        self.wire_ptrs = [int(p) for p in self.h_wire.buffer_ptrs] if self.h_wire else self.grad_ptrs
        self.mult_ptrs = [p + self.mult_off for p in self.sig_ptrs] if codec.needs_scale else []
        self.epoch = 1                                           # next free barrier epoch
        per_rank_vecs = self.n // 4 // max(ctx.world_size, 1)
        self.blocks = int(min(ops.allreduce_max_blocks(), max(8, per_rank_vecs // 1024)))
        # algorithm: "twoshot" (peer loads/stores), "nvls" (in-switch multimem reduction, fp32 wire),
        # "oneshot" (small buffers); "auto" = nvls when the fabric exposes multicast (fp32 wire), else
        # oneshot if tiny, else twoshot.
        try:
            self.multicast_ptr = int(self.h_grad.multicast_ptr or 0)
        except Exception:
            self.multicast_ptr = 0
        import os
        if algorithm == "auto":
            # Measured (profiles/allreduce_sweep_{2,4,8}gpu.txt): with 4 or more ranks the in-switch reduction
            # wins at every size; between two GPUs the peer load/store kernel does (622 vs 399 GB/s bus at 256 MB:
            # the switch has nothing to save with one peer).  B200SEG_NVLS=0 forces the peer kernels.
            algorithm = "nvls" if (os.environ.get("B200SEG_NVLS", "1") != "0" and self.multicast_ptr
                                   and codec.name == "fp32" and ctx.world_size > 2) else "twoshot"
        if algorithm == "nvls" and not (self.multicast_ptr and codec.name == "fp32"):
            raise RuntimeError("NVLS all-reduce needs multicast support and the fp32 codec")
        self.algorithm = algorithm
        if algorithm == "nvls":
            # multimem kernels saturate the fabric with few CTAs (4 x 16 B in flight per thread); more CTAs only
            # add contention: 32 CTAs 810 / 708 / 481 GB/s bus at 256 / 64 / 16 MB on 8 GPUs against 748 / 561 / 384 with 256
            self.blocks = int(os.environ.get("B200SEG_NVLS_BLOCKS", "32"))
        # lossy codecs: quantize-on-send push kernel (B200SEG_PUSH=0: the older encode pass + pull two-shot)
        self.push = codec.name != "fp32" and algorithm == "twoshot" and os.environ.get("B200SEG_PUSH", "1") != "0"
        self.oneshot_out: Optional[torch.Tensor] = None
        if (self.n * 4 <= ONESHOT_MAX_BYTES and self.algorithm != "nvls" and not self.push) or self.algorithm == "oneshot":
            self.oneshot_out = torch.empty(self.n, dtype=torch.float32, device=dev)
        # Fused mode: the first `fused_bulk_elems` floats (the conv / convT weight gradients) are reduced
        # by the wgrad kernels themselves -- their epilogues add through the multicast alias, i.e. into
        # every replica at once (csrc/common.cuh grad_red_add).  Only the small tail goes through an
        # all-reduce kernel, whose entry barrier is also what orders those remote adds before the
        # optimizer reads the buffer.
        self.fused_bulk = 0
        if fused_bulk_elems:
            if not (self.multicast_ptr and codec.name == "fp32"):
                raise RuntimeError("fused wgrad+all-reduce needs NVSwitch multicast and the fp32 codec")
            if fused_bulk_elems % 64 or fused_bulk_elems > n:
                raise ValueError("fused_bulk_elems must be a multiple of 64 within the buffer")
            self.fused_bulk = int(fused_bulk_elems)
            ops.set_grad_multicast(self.buffer_full, self.fused_bulk, self.multicast_ptr)
            self.tail_n = self.n - self.fused_bulk
        torch.cuda.synchronize(dev)
        dist.barrier(device_ids=[dev.index])

    # -- exchange overlapped with backward (csrc/comm/overlap.cu) ------------------------------------
    def setup_overlap(self, buckets, timeout_s: float = 30.0, blocks: int = 0) -> None:
        """Allocate the bucket signal pad (symmetric), the local exchange state and one ``GradArm`` per
        bucket.  ``buckets``: ``parallel.overlap.plan_buckets`` output for THIS buffer."""
        import os
        import torch.distributed._symmetric_memory as symm_mem
        ops = torch.ops.b200seg
        if self.codec.name != "fp32":
            raise RuntimeError("the overlapped exchange reduces fp32 gradients in place; use codec fp32")
        if len(buckets) > ops.overlap_max_buckets():
            raise ValueError("too many buckets")
        dev = self.buffer.device
        world, rank = self.ctx.world_size, self.ctx.rank
        self.ovl_pad = symm_mem.empty(int(ops.overlap_pad_words()), dtype=torch.int32, device=dev)
        self.ovl_pad.zero_()
        self.h_ovl = symm_mem.rendezvous(self.ovl_pad, self.group)
        self.ovl_pad_ptrs = [int(p) for p in self.h_ovl.buffer_ptrs]
        # local state: [0] exchange number, [1] CTA counter of the final kernel, [2 + b] arrival counter of bucket b
        self.ovl_state = torch.zeros(2 + len(buckets), dtype=torch.int32, device=dev)
        self.ovl_state[0] = 1
        base = self.ovl_state.data_ptr()
        arms = torch.zeros(len(buckets), 19, dtype=torch.int64)
        for b in buckets:
            arms[b.index, 0] = base + 4 * (2 + b.index)
            arms[b.index, 1] = base
            for r in range(world):
                arms[b.index, 2 + r] = self.ovl_pad_ptrs[r] + 4 * int(ops.overlap_flag_index(b.index, rank))
            arms[b.index, 18] = world                      # int world + 4 bytes of padding (little endian)
        self.ovl_arms = arms.to(dev)
        self.ovl_buckets = list(buckets)
        self.ovl_timeout_s = float(timeout_s)
        self.ovl_blocks = int(blocks or os.environ.get("B200SEG_OVL_BLOCKS", "16"))
        self.ovl_abort_index = int(ops.overlap_abort_index())
        torch.cuda.synchronize(dev)
        dist.barrier(device_ids=[dev.index])

    def arm_ptr(self, bucket_index: int) -> int:
        return self.ovl_arms.data_ptr() + 19 * 8 * bucket_index

    @torch.no_grad()
    def reduce_bucket(self, bucket, self_publish: bool) -> None:
        """Queue the reduction of one bucket on the CURRENT stream (the engine runs it on its side stream)."""
        from ..ops import functional as Fn
        offs = [o for o, _ in bucket.ranges]
        cnts = [n for _, n in bucket.ranges]
        torch.ops.b200seg.bucket_reduce(self.multicast_ptr, [] if self.multicast_ptr else self.grad_ptrs,
                                        self.ovl_pad_ptrs, self.ovl_state, self.ctx.rank, bucket.index,
                                        bool(self_publish or bucket.final), bool(bucket.final), offs, cnts, 1.0,
                                        self.ovl_timeout_s, self.ovl_blocks)
        Fn._tick()

    def check_abort(self) -> None:
        """Raise ``PeerLost`` if a bounded wait on the device gave up (synchronizes the device)."""
        torch.cuda.synchronize(self.signal.device)
        word = int(self.signal[self.abort_word].item()) & 0xFFFFFFFF
        pad = getattr(self, "ovl_pad", None)
        if not word and pad is not None:
            word = int(pad[self.ovl_abort_index].item()) & 0xFFFFFFFF
        if word:
            from .star import PeerLost
            raise PeerLost(f"gradient exchange timed out on rank {word & 0xFFFF} after {self.timeout_s:.0f} s "
                           "(a peer did not arrive); the step's gradients are not reduced")

    @property
    def multicast_supported(self) -> bool:
        try:
            return bool(self.h_grad.has_multicast_support)
        except Exception:
            return False

    @torch.no_grad()
    def all_reduce_(self, post_scale: float = 1.0) -> torch.Tensor:
        """Sum over ranks (times ``post_scale``) into ``buffer`` on every rank."""
        from ..ops import functional as Fn
        ops = torch.ops.b200seg
        c = self.codec
        epoch = self.epoch
        self.epoch += 2
        if c.name != "fp32" and self.push:
            # quantize-on-send: the all-reduce kernel encodes in registers and pushes into the owners' inboxes;
            # a lossy codec costs one extra launch (the global max-abs), not an encode pass
            scale = Fn.codec_amax(self.buffer_full) if c.needs_scale else None
            ops.allreduce_push(self.wire_ptrs, self.grad_ptrs, self.sig_ptrs,
                               [p + self.mult_off for p in self.sig_ptrs], scale, c.name, self.ctx.rank, self.n,
                               float(post_scale), epoch, self.blocks)
            Fn._tick()
            return self.buffer
        if c.name != "fp32":
            scale = Fn.codec_amax(self.buffer_full) if c.needs_scale else torch.ones((), device=self.buffer.device)
            ops.codec_encode(self.buffer_full, scale, self.wire, c.name)
            Fn._tick()
            if c.needs_scale:
                levels = c.levels if c.levels is not None else 448.0
                ops.write_mult(scale, float(levels), self.mult_slot)
                Fn._tick()
        if self.fused_bulk:
            if self.tail_n:
                off = self.fused_bulk * 4
                ops.allreduce_nvls(self.sig_ptrs, self.multicast_ptr + off, self.ctx.rank, self.tail_n,
                                   float(post_scale), epoch, 8)
            else:
                ops.peer_barrier(self.sig_ptrs, self.ctx.rank, epoch)
            Fn._tick()
            if post_scale != 1.0:
                self.buffer_full[:self.fused_bulk].mul_(post_scale)
            return self.buffer
        if self.algorithm == "nvls":
            ops.allreduce_nvls(self.sig_ptrs, self.multicast_ptr, self.ctx.rank, self.n, float(post_scale), epoch,
                               self.blocks)
            Fn._tick()
            return self.buffer
        ops.allreduce(self.wire_ptrs, self.grad_ptrs, self.sig_ptrs, self.mult_ptrs, self.ctx.rank, self.n,
                      self.wire_kind, float(post_scale), epoch, self.oneshot_out, self.blocks)
        Fn._tick()
        if self.oneshot_out is not None:
            self.buffer_full.copy_(self.oneshot_out)
        return self.buffer

    @torch.no_grad()
    def barrier(self) -> None:
        """Cross-GPU barrier on the launching stream (one epoch).  Fused mode calls it after the
        optimizer step: no rank may start adding into a peer's gradient buffer before that peer has
        consumed and zeroed it."""
        from ..ops import functional as Fn
        torch.ops.b200seg.peer_barrier(self.sig_ptrs, self.ctx.rank, self.epoch)
        self.epoch += 1
        Fn._tick()

    def suspend_fused(self) -> None:
        """Fused mode: stop the wgrad launchers from adding through the multicast alias (graph warm-up)."""
        if self.fused_bulk:
            torch.ops.b200seg.set_grad_multicast(self.buffer_full, 0, 0)

    def resume_fused(self) -> None:
        if self.fused_bulk:
            torch.ops.b200seg.set_grad_multicast(self.buffer_full, self.fused_bulk, self.multicast_ptr)

    def close(self) -> None:
        if self.fused_bulk:
            torch.ops.b200seg.set_grad_multicast(self.buffer_full, 0, 0)
            self.fused_bulk = 0
