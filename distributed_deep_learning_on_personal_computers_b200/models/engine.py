"""Static-shape training engine for the U-Net: explicit forward, loss and backward.

The reference's training step is ``model(x) -> CrossEntropyLoss -> loss.backward()`` through
eager autograd (``R:754-756``), one library kernel per module.  Here the step is a fixed plan
over pre-allocated NHWC buffers:

* no autograd tape and no allocator traffic in the hot loop, so the whole micro-batch (and the
  optimizer step) is capturable in one CUDA graph;
* ``torch.cat`` (``R:616``) does not exist: the encoder's BN+ReLU kernel and the transposed-conv
  epilogue write into channel slices of one concat buffer per decoder level;
* BN statistics come out of the conv epilogue, BN-apply+ReLU(+2x2 max-pool) is one pass, the
  1x1 head + softmax cross-entropy + accuracy + the head's whole backward is one pass;
* weight gradients are accumulated in fp32 straight into the flat gradient buffer that the
  all-reduce and the fused optimizer work on (``models.unet.UNet.flatten_parameters``).

Every op goes through ``ops.functional`` which runs the sm_100a kernel on a B200 and the
plain-PyTorch oracle (``ops.reference``) on CPU, so this file is testable against autograd of
``models.unet.UNet.forward`` on any machine.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from ..ops import functional as Fn
from .unet import UNet, ConvSpec

IN_PAD = 16   # the 3 input channels are zero-padded to one UMMA K-step (16 bf16)


@dataclass
class _Stage:
    """One conv3x3+BN+ReLU stage and the tensors it owns."""
    spec: ConvSpec
    x: torch.Tensor = None            # input activation (N,H,W,Cin) (maybe a concat buffer)
    y: torch.Tensor = None            # raw conv output (N,H,W,Cout)
    out: torch.Tensor = None          # activated output (maybe a slice of a concat buffer)
    pool: Optional[torch.Tensor] = None
    ss: torch.Tensor = None           # (4,C) scale, shift, mean, invstd of this micro-batch
    stats: torch.Tensor = None
    dy: torch.Tensor = None           # gradient w.r.t. y
    d_out: Optional[torch.Tensor] = None   # gradient w.r.t. out   (set by the consumer)
    d_pool: Optional[torch.Tensor] = None  # gradient w.r.t. pool  (set by the consumer)
    dx: Optional[torch.Tensor] = None      # where to write the data gradient (None = skip)


class NativeEngine:
    """Runs forward + loss + backward of one micro-batch over the model's flat buffers."""

    def __init__(self, model: UNet, batch: int, height: int, width: int,
                 act_dtype: Optional[torch.dtype] = None, conv_impl: str = "auto",
                 fused_tail: Optional[bool] = None):
        if height % 32 or width % 32:
            raise ValueError("H and W must be multiples of 32 (five 2x2 pools, R:644-648)")
        self.model = model
        flat = model.flat
        self.flat_p, self.flat_g, self.offsets = flat["param"], flat["grad"], flat["offsets"]
        self.device = self.flat_p.device
        self.on_gpu = self.device.type == "cuda"
        self.act_dtype = act_dtype or (torch.bfloat16 if self.on_gpu else torch.float32)
        self.conv_impl = conv_impl
        self.N, self.H, self.W = batch, height, width
        self.classes = model.out_classes
        # One pass over the last conv's raw output does its BN+ReLU, the head with loss and backward, and the
        # two sums of that BN's backward (csrc/elem/head_tail.cu); the last activation is never materialised.
        if fused_tail is None:
            fused_tail = os.environ.get("B200SEG_HEAD_TAIL", "1") != "0"
        self._want_fused_tail = bool(fused_tail)
        self.bilinear = model.up_sample_mode == "bilinear"
        # low-precision copy of the weights that the tensor-core kernels read
        if self.act_dtype == torch.float32:
            self.flat_lp = self.flat_p
        else:
            self.flat_lp = self.flat_p.to(self.act_dtype)
        self._params = dict(model.named_parameters())
        self._buffers = dict(model.named_buffers())
        if self.on_gpu:
            from ..ops import loader
            loader.arm_error_record()            # a timed-out barrier wait leaves its address behind before it traps
        self._build()
        self.training = True
        # The stem reads 3 real + 13 zero channels (one UMMA K step).  Its bf16 weights are kept
        # zero-padded to (Cout, 3, 3, IN_PAD) next to the flat buffer, refreshed every forward, and its
        # weight gradient is produced padded and folded back into the flat gradient.
        self.stem = "down_conv1.double_conv.double_conv.0"
        cout = self._params[self.stem + ".weight"].shape[0]
        self.stem_pad = self.on_gpu and model.in_channels < IN_PAD
        if self.stem_pad:
            self.stem_w = torch.zeros(cout, 3, 3, IN_PAD, dtype=self.act_dtype, device=self.device)
            self.stem_dw = torch.zeros(cout, 3, 3, IN_PAD, dtype=torch.float32, device=self.device)
        # gradient exchange overlapped with backward (set per call by train_microbatch(overlap=...))
        self._ov = None
        self._ov_trigger: Dict[str, object] = {}
        self._comm_stream = None
        self._ov_events: List[object] = []

    # -- helpers -------------------------------------------------------------------------------
    def _new(self, *shape, dtype=None):
        return torch.zeros(*shape, dtype=dtype or self.act_dtype, device=self.device)

    def _seg(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        off, n = self.offsets[name]
        return flat[off:off + n]

    def _w3(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        p = self._params[name]
        co, ci = p.shape[0], p.shape[1]
        return self._seg(flat, name).view(co, 3, 3, ci)

    @torch.no_grad()
    def refresh_lowp(self) -> None:
        """Re-cast the weights after an optimizer step that did not do it itself."""
        if self.flat_lp is not self.flat_p:
            self.flat_lp.copy_(self.flat_p)

    # -- plan ------------------------------------------------------------------------------------
    def _build(self) -> None:
        m, N, H, W = self.model, self.N, self.H, self.W
        specs = m.conv_specs()
        b = m.base_channels
        wd = [b, 2 * b, 4 * b, 8 * b, 8 * b]
        c_up = {5: wd[4], 4: wd[3], 3: wd[3], 2: wd[2], 1: wd[1]}
        c_skip = {i: wd[i - 1] for i in range(1, 6)}
        self.c_up, self.c_skip = c_up, c_skip
        res = lambda lvl: (H >> lvl, W >> lvl)

        self.x_in = self._new(N, H, W, IN_PAD)
        # concat buffers, one per decoder level i: [up (c_up) | skip (c_skip)]  (R:616 order)
        self.cat = {i: self._new(N, *res(i - 1), c_up[i] + c_skip[i]) for i in range(1, 6)}
        self.d_cat = {i: self._new(N, *res(i - 1), c_up[i] + c_skip[i]) for i in range(1, 6)}

        self.stages: Dict[str, Tuple[_Stage, _Stage]] = {}
        order: List[_Stage] = []
        x = self.x_in
        d_x = None
        for i in range(1, 6):                                   # encoder
            s0, s1 = (_Stage(sp) for sp in specs[f"down_conv{i}"])
            h, w = res(i - 1)
            c = wd[i - 1]
            s0.x, s0.y, s0.out = x, self._new(N, h, w, c), self._new(N, h, w, c)
            s0.dy, s0.dx = self._new(N, h, w, c), d_x
            s1.x, s1.y = s0.out, self._new(N, h, w, c)
            s1.out = self.cat[i][..., c_up[i]:]
            s1.pool = self._new(N, h // 2, w // 2, c)
            s1.dy = self._new(N, h, w, c)
            s1.dx = s0.d_out = self._new(N, h, w, c)
            s1.d_out = self.d_cat[i][..., c_up[i]:]
            s1.d_pool = self._new(N, h // 2, w // 2, c)
            x, d_x = s1.pool, s1.d_pool
            self.stages[f"down_conv{i}"] = (s0, s1)
            order += [s0, s1]
        s0, s1 = (_Stage(sp) for sp in specs["double_conv"])      # bottleneck
        h, w = res(5)
        c = wd[4]
        s0.x, s0.y, s0.out, s0.dy, s0.dx = x, self._new(N, h, w, c), self._new(N, h, w, c), \
            self._new(N, h, w, c), d_x
        s1.x, s1.y, s1.out, s1.dy = s0.out, self._new(N, h, w, c), self._new(N, h, w, c), \
            self._new(N, h, w, c)
        s1.dx = s0.d_out = self._new(N, h, w, c)
        s1.d_out = self._new(N, h, w, c)
        self.stages["double_conv"] = (s0, s1)
        order += [s0, s1]
        below, d_below = s1.out, s1.d_out
        self.up_io = {}
        for i in (5, 4, 3, 2, 1):                                # decoder
            s0, s1 = (_Stage(sp) for sp in specs[f"up_conv{i}"])
            h, w = res(i - 1)
            c = s0.spec.cout
            self.up_io[i] = (below, d_below)
            s0.x, s0.y, s0.out, s0.dy = self.cat[i], self._new(N, h, w, c), \
                self._new(N, h, w, c), self._new(N, h, w, c)
            s0.dx = self.d_cat[i]
            s1.x, s1.y, s1.out, s1.dy = s0.out, self._new(N, h, w, c), self._new(N, h, w, c), \
                self._new(N, h, w, c)
            s1.dx = s0.d_out = self._new(N, h, w, c)
            s1.d_out = self._new(N, h, w, c)
            below, d_below = s1.out, s1.d_out
            self.stages[f"up_conv{i}"] = (s0, s1)
            order += [s0, s1]
        self.order = order
        self.head_in, self.d_head_in = below, d_below
        self.tail = s1                                           # up_conv1's second stage feeds the head
        self.fused_tail = self._want_fused_tail and Fn.head_tail_supported(s1.y, self.classes)
        if self.fused_tail:
            self.tail_rows = Fn.head_tail_rows(s1.y)
            self.tail_partial = self._new(self.tail_rows, 2, s1.spec.cout, dtype=torch.float32)
        for st in order:
            st.ss = self._new(4, st.spec.cout, dtype=torch.float32)
        self.loss_acc = self._new(2, dtype=torch.float32)       # [loss sum, correct] of last call
        self.pred = torch.zeros(N, H, W, dtype=torch.uint8, device=self.device)
        self.labels = torch.zeros(N, H, W, dtype=torch.uint8, device=self.device)

    # -- one stage ---------------------------------------------------------------------------------
    def _stage_fwd(self, st: _Stage) -> None:
        sp = st.spec
        w = self._w3(self.flat_lp, sp.name + ".weight")
        if self.stem_pad and sp.name == self.stem:
            torch.ops.b200seg.pad_rows(w, self.stem_w)       # 3 -> 16 channels, zero padded (no ATen op in the graph)
            Fn._tick()
            w = self.stem_w
        bias = self._seg(self.flat_p, sp.name + ".bias")
        # the conv bias is NOT added by the conv: BatchNorm absorbs it (see ops.reference.bn_finalize)
        st.stats = Fn.conv3x3_fwd(st.x, w, None, st.y, want_stats=self.training,
                                  impl=self.conv_impl)
        bn = sp.bn
        Fn.bn_finalize(st.stats, st.y.shape[0] * st.y.shape[1] * st.y.shape[2],
                       self._seg(self.flat_p, bn + ".weight"), self._seg(self.flat_p, bn + ".bias"),
                       self._buffers[bn + ".running_mean"], self._buffers[bn + ".running_var"],
                       self._buffers[bn + ".num_batches_tracked"], st.ss, training=self.training,
                       conv_bias=bias)
        if self.fused_tail and st is self.tail:
            return                                   # the head kernel applies this BN+ReLU on load
        Fn.bn_relu_pool_fwd(st.y, st.ss, st.out, st.pool)

    def _stage_bwd(self, st: _Stage, need_dx: bool = True, colsum: bool = False):
        """``colsum``: also return the per-CTA column sums of ``dx`` from the dgrad epilogue."""
        sp = st.spec
        bn = sp.bn
        if self.fused_tail and st is self.tail:
            # the head kernel left sum g and sum g*y of this stage in tail_partial: finalize + apply only
            Fn.bn_relu_bwd_from_partial(self.tail_partial, self.tail_rows, st.d_out, st.y, st.ss,
                                        self._seg(self.flat_p, bn + ".weight"),
                                        self._seg(self.flat_g, bn + ".weight"),
                                        self._seg(self.flat_g, bn + ".bias"), st.dy)
        else:
            Fn.bn_relu_pool_bwd(st.d_out, st.d_pool if st.pool is not None else None, st.y, st.ss,
                                self._seg(self.flat_p, bn + ".weight"),
                                self._seg(self.flat_g, bn + ".weight"),
                                self._seg(self.flat_g, bn + ".bias"),
                                self._seg(self.flat_g, sp.name + ".bias"), st.dy)
        dw = self._w3(self.flat_g, sp.name + ".weight")
        if self.stem_pad and sp.name == self.stem:
            # stem_dw is zero on entry (allocated zeroed, re-zeroed by the fold below)
            Fn.conv3x3_wgrad(st.x, st.dy, self.stem_dw, impl=self.conv_impl)
            torch.ops.b200seg.unpad_add_zero(self.stem_dw, dw)
            Fn._tick()
        else:
            bucket = self._arm(sp.name + ".weight", sp.cin, sp.cout)
            Fn.conv3x3_wgrad(st.x, st.dy, dw, impl=self.conv_impl)
            self._bucket_done(bucket)
        if need_dx and st.dx is not None:
            w_lp = self._w3(self.flat_lp, sp.name + ".weight")
            if colsum:
                return Fn.conv3x3_dgrad_stats(st.dy, w_lp, st.dx, impl=self.conv_impl)
            Fn.conv3x3_dgrad(st.dy, w_lp, st.dx, impl=self.conv_impl)
        return None

    # -- public steps --------------------------------------------------------------------------------
    @torch.no_grad()
    def load_batch(self, images_u8: torch.Tensor, labels_u8: torch.Tensor) -> None:
        """Device-side input preparation (R:737-741): u8 HWC -> /255 -> padded NHWC; labels u8."""
        Fn.prep_input(images_u8, self.x_in)
        self.labels.copy_(labels_u8, non_blocking=True)

    @torch.no_grad()
    def forward(self) -> None:
        for i in range(1, 6):
            s0, s1 = self.stages[f"down_conv{i}"]
            self._stage_fwd(s0)
            self._stage_fwd(s1)
        s0, s1 = self.stages["double_conv"]
        self._stage_fwd(s0)
        self._stage_fwd(s1)
        for i in (5, 4, 3, 2, 1):
            below, _ = self.up_io[i]
            up_slice = self.cat[i][..., :self.c_up[i]]
            if self.bilinear:
                Fn.upsample_bilinear2x_fwd(below, up_slice)
            else:
                name = f"up_conv{i}.up_sample"
                Fn.convT2x2_fwd(below, self._params[name + ".weight"],
                                self._convT_lp(name), self._seg(self.flat_p, name + ".bias"),
                                up_slice)
            s0, s1 = self.stages[f"up_conv{i}"]
            self._stage_fwd(s0)
            self._stage_fwd(s1)

    def _convT_lp(self, name: str) -> torch.Tensor:
        p = self._params[name + ".weight"]
        ci, co = p.shape[0], p.shape[1]
        return self._seg(self.flat_lp, name + ".weight").view(4 * co, ci)

    def _head(self, with_grad: bool, grad_scale: float) -> None:
        wh = self._seg(self.flat_p, "conv_last.weight").view(self.classes, -1)
        bh = self._seg(self.flat_p, "conv_last.bias")
        if self.fused_tail:
            st = self.tail
            if with_grad:
                dwh = self._seg(self.flat_g, "conv_last.weight").view(self.classes, -1)
                dbh = self._seg(self.flat_g, "conv_last.bias")
                Fn.head_tail_fwd_bwd(st.y, st.ss, wh, bh, self.labels, grad_scale, self.d_head_in, dwh, dbh,
                                     self.pred, self.loss_acc, self.tail_partial)
            else:
                Fn.head_tail_fwd_bwd(st.y, st.ss, wh, bh, self.labels, 0.0, None, None, None, self.pred,
                                     self.loss_acc, None)
            return
        if with_grad:
            dwh = self._seg(self.flat_g, "conv_last.weight").view(self.classes, -1)
            dbh = self._seg(self.flat_g, "conv_last.bias")
            Fn.head_ce_fwd_bwd(self.head_in, wh, bh, self.labels, grad_scale, self.d_head_in,
                               dwh, dbh, self.pred, self.loss_acc)
        else:
            Fn.head_ce_fwd_bwd(self.head_in, wh, bh, self.labels, 0.0, None, None, None,
                               self.pred, self.loss_acc)

    # -- exchange under backward (parallel/overlap.py, csrc/comm/overlap.cu) ----------------------------
    def _begin_overlap(self, ov) -> None:
        """``ov``: ``parallel.symm.SymmetricAllReduce`` after ``setup_overlap`` (or None)."""
        self._ov = ov
        if ov is None:
            return
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(self.device)
            self._ov_events = [torch.cuda.Event() for _ in range(len(ov.ovl_buckets) + 1)]
        self._ov_trigger = {b.trigger: b for b in ov.ovl_buckets if b.trigger is not None}
        self._ov_armed = (Fn._conv_impl(self.conv_impl) != "cudnn" and
                          os.environ.get("B200SEG_WGRAD", "tcgen05") != "cudnn")

    def _arm(self, param_name: str, cin: int, cout: int):
        """Before the wgrad call that completes a bucket: make its last kernel the all-reduce trigger."""
        if self._ov is None:
            return None
        b = self._ov_trigger.get(param_name)
        if b is None:
            return None
        armed = self._ov_armed and Fn._use_tc(cin, cout, self.conv_impl)
        if armed:
            torch.ops.b200seg.arm_next_wgrad(self._ov.arm_ptr(b.index))
        return (b, armed)

    def _bucket_done(self, armed_bucket) -> None:
        if armed_bucket is None:
            return
        b, armed = armed_bucket
        if armed:
            torch.ops.b200seg.arm_next_wgrad(0)             # never leave a stale trigger behind
        self._queue_reduce(b, self_publish=not armed)

    def _queue_reduce(self, b, self_publish: bool) -> None:
        cur = torch.cuda.current_stream(self.device)
        ev = self._ov_events[b.index]
        ev.record(cur)                                      # the bucket's producers are all queued before this
        self._comm_stream.wait_event(ev)
        with torch.cuda.stream(self._comm_stream):
            self._ov.reduce_bucket(b, self_publish)

    def _end_overlap(self) -> None:
        """The final bucket (first encoder block, BN / bias / stem / head gradients), then join."""
        ov = self._ov
        if ov is None:
            return
        self._queue_reduce(ov.ovl_buckets[-1], self_publish=True)
        done = self._ov_events[-1]
        done.record(self._comm_stream)
        torch.cuda.current_stream(self.device).wait_event(done)
        self._ov = None

    @torch.no_grad()
    def backward(self) -> None:
        for i in (1, 2, 3, 4, 5):
            s0, s1 = self.stages[f"up_conv{i}"]
            self._stage_bwd(s1)
            # the up-sampled slice of d_cat is written by this dgrad; its column sums are the transposed
            # conv's bias gradient
            part = self._stage_bwd(s0, colsum=not self.bilinear)
            below, d_below = self.up_io[i]
            d_up = self.d_cat[i][..., :self.c_up[i]]
            if self.bilinear:
                Fn.upsample_bilinear2x_bwd(d_up, d_below)
            else:
                name = f"up_conv{i}.up_sample"
                bucket = self._arm(name + ".weight", self.c_up[i], self.c_up[i])
                Fn.convT2x2_bwd(below, d_up, self._params[name + ".weight"], self._convT_lp(name),
                                self._params[name + ".weight"].grad,
                                self._seg(self.flat_g, name + ".bias"), d_below, d_up_partials=part)
                self._bucket_done(bucket)
        s0, s1 = self.stages["double_conv"]
        self._stage_bwd(s1)
        self._stage_bwd(s0)
        for i in (5, 4, 3, 2, 1):
            s0, s1 = self.stages[f"down_conv{i}"]
            self._stage_bwd(s1)
            self._stage_bwd(s0, need_dx=(i != 1))

    @torch.no_grad()
    def train_microbatch(self, loss_scale: Optional[float] = None, overlap=None) -> torch.Tensor:
        """Forward + loss + backward on the batch already in ``x_in`` / ``labels``.

        Gradients are ACCUMULATED into the flat gradient buffer (R:756 semantics).  Returns the
        device tensor ``[loss_sum, correct]`` (read it back only when logging).

        ``overlap``: the symmetric-memory reducer; when given, this micro-batch ends an accumulation
        window and the gradient buckets are all-reduced across ranks while backward runs (each
        bucket's reduction is triggered by the wgrad kernel that completes it).  On return the
        flat gradient buffer holds the SUM over ranks, bit-identical everywhere (R:402-433).
        """
        self.training = True
        pixels = self.N * self.H * self.W
        self.forward()
        self._head(True, (1.0 / pixels) if loss_scale is None else loss_scale)
        self._begin_overlap(overlap)
        self.backward()
        self._end_overlap()
        return self.loss_acc

    @torch.no_grad()
    def evaluate(self) -> torch.Tensor:
        """Forward only with running statistics (the reference never does this)."""
        self.training = False
        self.forward()
        self._head(False, 0.0)
        self.training = True
        return self.loss_acc

    def logits_free_prediction(self) -> torch.Tensor:
        return self.pred
