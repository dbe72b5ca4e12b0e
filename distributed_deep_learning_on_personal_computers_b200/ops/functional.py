"""Dispatch between the sm_100a kernels (``csrc/``, loaded from the in-tree shared library) and
the plain-PyTorch oracles (``ops.reference``).

Rule: a CUDA tensor ALWAYS goes to the native kernel.  If the shared library is missing on a
GPU box the call raises instead of silently running a PyTorch fallback (set
``B200SEG_ALLOW_FALLBACK=1`` to override, e.g. to produce oracle values on the GPU in a test).
CPU tensors run the oracle.  The two conv implementations selectable with ``impl`` are
``"tcgen05"`` (ours) and ``"cudnn"`` (library; used only as a comparison arm and oracle).

``launch_count()`` returns how many of OUR kernels were launched since the last reset; bench.py
reports it as ``gpu_launches``.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import reference as ref
from . import loader

_LAUNCHES = 0


def _tick(n: int = 1) -> None:
    global _LAUNCHES
    _LAUNCHES += n


def launch_count() -> int:
    return _LAUNCHES


def reset_launch_count() -> None:
    global _LAUNCHES
    _LAUNCHES = 0


def native_available() -> bool:
    return loader.library() is not None


def _native(t: torch.Tensor) -> bool:
    """True -> run the sm_100a kernel.  Raises on a GPU tensor without the library."""
    if not t.is_cuda:
        return False
    if loader.library() is not None:
        return True
    if os.environ.get("B200SEG_ALLOW_FALLBACK") == "1":
        return False
    raise RuntimeError(
        "CUDA tensor passed to a b200seg op but the native library is not built/loaded "
        f"({loader.load_error()}). Run `python -c 'import __graft_entry__ as g; g.build()'`.")


def _ops():
    return torch.ops.b200seg


def _conv_impl(impl: Optional[str]) -> str:
    impl = impl or "auto"
    if impl == "auto":
        impl = os.environ.get("B200SEG_CONV", "tcgen05")
    return impl


# ---------------------------------------------------------------------------------------------
def _use_tc(cin: int, cout: int, impl) -> bool:
    """tcgen05 path unless the caller asked for the library arm or the shape is unsupported."""
    if _conv_impl(impl) == "cudnn":
        return False
    return bool(_ops().conv_supported(int(cin), int(cout)))


def conv3x3_fwd(x, w, bias, y, want_stats=True, impl=None):
    """y = conv3x3(x, w) + bias; returns the BN statistics handle (see ``bn_finalize``)."""
    if _native(x):
        cout, cin = w.shape[0], w.shape[3]
        if _use_tc(cin, cout, impl):
            part = _ops().conv3x3_fwd(x, w, bias, y, bool(want_stats))
            _tick()
            return part if want_stats else None
        # library comparison arm (cuDNN through torch), statistics by our reduction kernel
        yy = torch.nn.functional.conv2d(x[..., :cin].permute(0, 3, 1, 2), w.permute(0, 3, 1, 2),
                                        None if bias is None else bias.to(x.dtype), padding=1)
        y.copy_(yy.permute(0, 2, 3, 1))
        if not want_stats:
            return None
        part = torch.empty(_ops().stats_rows(), 2, y.shape[-1], dtype=torch.float32, device=y.device)
        _ops().channel_stats(y, part)
        _tick()
        return part
    yy, stats = ref.conv3x3_fwd(x, w, bias, want_stats)
    y.copy_(yy)
    return stats


def conv3x3_dgrad(dy, w, dx, impl=None):
    if _native(dy):
        if _use_tc(w.shape[3], w.shape[0], impl):
            _ops().conv3x3_dgrad(dy, w, dx)
            _tick()
            return dx
        g = torch.nn.functional.conv_transpose2d(dy.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), padding=1)
        dx.copy_(g.permute(0, 2, 3, 1))
        return dx
    return ref.conv3x3_dgrad(dy, w, dx)


def conv3x3_dgrad_stats(dy, w, dx, impl=None):
    """``conv3x3_dgrad`` that also returns per-CTA column sums of ``dx`` (``[rows, 2, Cin]``: sum and sum
    of squares) from the kernel's epilogue, or ``None`` where that kernel is not in use.  The transposed
    conv's bias gradient is such a column sum (``convT2x2_bwd(..., d_up_partials=...)``)."""
    if _native(dy) and _use_tc(w.shape[3], w.shape[0], impl):
        part = _ops().conv3x3_dgrad_stats(dy, w, dx)
        _tick()
        return part
    conv3x3_dgrad(dy, w, dx, impl=impl)
    return None


def conv3x3_wgrad(x, dy, dw, impl=None):
    if _native(x):
        cout, _, _, cin = dw.shape
        if _use_tc(cin, cout, impl) and os.environ.get("B200SEG_WGRAD", "tcgen05") != "cudnn":
            _ops().conv3x3_wgrad(x, dy, dw)
            _tick()
            return dw
        g = torch.nn.grad.conv2d_weight(x[..., :cin].permute(0, 3, 1, 2), (cout, cin, 3, 3),
                                        dy.permute(0, 3, 1, 2), padding=1)
        dw.add_(g.permute(0, 2, 3, 1))
        return dw
    return ref.conv3x3_wgrad(x, dy, dw)


def bn_finalize(stats, count, gamma, beta, running_mean, running_var, num_batches_tracked, ss,
                momentum=0.1, eps=1e-5, training=True, conv_bias=None):
    if _native(ss):
        _ops().bn_finalize(stats, int(count), gamma, beta, conv_bias, running_mean, running_var,
                           num_batches_tracked, ss, float(momentum), float(eps), bool(training))
        _tick()
        return ss
    out = ref.bn_finalize(stats, count, gamma, beta, running_mean, running_var, momentum, eps,
                          training, conv_bias)
    if training:
        num_batches_tracked.add_(1)
    ss.copy_(out)
    return ss


def bn_relu_pool_fwd(y, ss, out, pool=None):
    if _native(y):
        _ops().bn_relu_pool_fwd(y, ss, out, pool)
        _tick()
        return
    ref.bn_relu_pool_fwd(y, ss, out, pool)


def bn_relu_pool_bwd(d_out, d_pool, y, ss, gamma, dgamma, dbeta, dbias, dy):
    if _native(y):
        _ops().bn_relu_pool_bwd(d_out, d_pool, y, ss, gamma, dgamma, dbeta, dbias, dy)
        _tick(2)
        return
    ref.bn_relu_pool_bwd(d_out, d_pool, y, ss, gamma, dgamma, dbeta, dbias, dy)


def convT2x2_fwd(x, w_param, w_lp, bias, out, impl=None):
    """``w_param``: logical (Cin,Cout,2,2) fp32 parameter; ``w_lp``: (4*Cout, Cin) kernel copy."""
    if _native(x):
        cout4, cin = w_lp.shape
        if _use_tc(cin, cout4 // 4, impl):
            _ops().convT2x2_fwd(x, w_lp, bias, out)
            _tick()
            return
        wl = w_lp.view(2, 2, cout4 // 4, cin).permute(3, 2, 0, 1)
        up = torch.nn.functional.conv_transpose2d(x.permute(0, 3, 1, 2), wl, bias.to(x.dtype), stride=2)
        out.copy_(up.permute(0, 2, 3, 1))
        return
    ref.convT2x2_fwd(x, w_param, bias, out)


def convT2x2_bwd(x, d_up, w_param, w_lp, dw_param, dbias, dx, impl=None, d_up_partials=None):
    """Backward of the transposed conv.  Data gradient = a stride-2 2x2 conv of d_up (a tap GEMM over
    the four quadrants, gathered by a 5-D TMA box); weight gradient = the wgrad kernel with the
    quadrant as the N tile.  ``dw_param`` is the logical (Cin,Cout,2,2) view of the flat gradient."""
    if _native(x):
        cout4, cin = w_lp.shape
        cout = cout4 // 4
        if _use_tc(cin, cout, impl):
            if d_up_partials is not None:        # column sums of d_up already came out of the dgrad that wrote it
                _ops().rows_sum_add(d_up_partials, cout, dbias)
                _tick()
            else:
                part = torch.empty(_ops().stats_rows(), 2, cout, dtype=torch.float32, device=x.device)
                _ops().channel_stats(d_up, part)
                _ops().rows_sum_add(part, cout, dbias)
                _tick(2)
            _ops().convT2x2_bwd(x, d_up, w_lp, _flat_of(dw_param).reshape(4 * cout, cin), dx)
            return
        wl = w_lp.view(2, 2, cout, cin).permute(3, 2, 0, 1)               # logical (Cin,Cout,2,2)
        g = d_up.permute(0, 3, 1, 2)
        dbias.add_(g.sum(dim=(0, 2, 3), dtype=torch.float32))
        dwl = torch.nn.grad.conv2d_weight(g, (cin, cout, 2, 2), x.permute(0, 3, 1, 2), stride=2)
        dw_param.add_(dwl)
        if dx is not None:
            dx.copy_(torch.nn.functional.conv2d(g, wl, stride=2).permute(0, 2, 3, 1))
        return
    ref.convT2x2_bwd(x, d_up, w_param, dw_param, dbias, dx)


def _flat_of(p_view: torch.Tensor) -> torch.Tensor:
    """Memory-order (dy, dx, Cout, Cin) view of a transposed-conv parameter view."""
    return p_view.permute(2, 3, 1, 0)


def upsample_bilinear2x_fwd(x, out):
    if _native(x):
        _ops().upsample_bilinear2x_fwd(x, out)
        _tick()
        return
    ref.upsample_bilinear2x_fwd(x, out)


def upsample_bilinear2x_bwd(d_up, dx):
    if _native(d_up):
        _ops().upsample_bilinear2x_bwd(d_up, dx)
        _tick()
        return
    ref.upsample_bilinear2x_bwd(d_up, dx)


def head_ce_fwd_bwd(a, w, b, labels, grad_scale, da, dw, db, pred, loss_acc):
    if _native(a):
        _ops().head_ce_fwd_bwd(a, w, b, labels, float(grad_scale), da, dw, db, pred, loss_acc)
        _tick()
        return loss_acc
    loss_acc.copy_(ref.head_ce_fwd_bwd(a, w, b, labels, grad_scale, da, dw, db, pred))
    return loss_acc


_PLAN_OPS = {"fprop": 0, "dgrad": 1, "wgrad": 2, "convT_fwd": 3, "convT_dgrad": 4}
_PLAN_FIELDS = ("kernel", "block_k", "block_n", "n_tiles", "ocb", "rows_per_unit", "weights_resident", "private_epilogue",
                "staging_bufs", "strip_slots", "stages", "grid", "cluster", "m_tiles", "smem_bytes")
_PLAN_KERNELS = {0: "tap_gemm/box", 1: "tap_gemm/strip", 2: "wgrad/split-k", 3: "wgrad/strip128", 4: "wgrad/row-pair"}


def conv_plan(op: str, n: int, h: int, w: int, cin: int, cout: int) -> dict:
    """Which tcgen05 kernel variant and tiling a conv of this shape takes (``B200ConvPlan`` in include/b200seg.h).
    Host-only: works without a GPU once the library is built; honours the ``B200SEG_*`` switches in effect."""
    from . import loader
    if loader.library() is None:
        raise RuntimeError(f"native library not available: {loader.load_error()}")
    vals = list(_ops().conv_plan(_PLAN_OPS[op], n, h, w, cin, cout))
    plan = dict(zip(_PLAN_FIELDS, vals))
    plan["kernel"] = _PLAN_KERNELS[plan["kernel"]]
    return plan


def head_tail_supported(y, classes: int) -> bool:
    """Can the last stage's BN+ReLU, the head and that BN's backward sums run as one pass over ``y``?"""
    if _native(y):
        return bool(_ops().head_tail_supported(int(y.shape[-1]), int(classes)))
    return True


def head_tail_rows(y) -> int:
    if _native(y):
        return int(_ops().head_tail_rows(y.shape[0] * y.shape[1] * y.shape[2], y.shape[3]))
    return 1


def head_tail_fwd_bwd(y, ss, w, b, labels, grad_scale, g_out, dw, db, pred, loss_acc, bn_partial):
    """BN(train)+ReLU on load + 1x1 head + CE + accuracy + head backward + BN-backward sums (head_tail.cu)."""
    if _native(y):
        _ops().head_tail_fwd_bwd(y, ss, w, b, labels, float(grad_scale), g_out, dw, db, pred, loss_acc, bn_partial)
        _tick()
        return loss_acc
    loss_acc.copy_(ref.head_tail_fwd_bwd(y, ss, w, b, labels, grad_scale, g_out, dw, db, pred, bn_partial))
    return loss_acc


def bn_relu_bwd_from_partial(partial, rows, d_out, y, ss, gamma, dgamma, dbeta, dy):
    if _native(y):
        _ops().bn_relu_bwd_from_partial(partial, int(rows), d_out, y, ss, gamma, dgamma, dbeta, dy)
        _tick(2)
        return
    ref.bn_relu_bwd_from_partial(partial, rows, d_out, y, ss, gamma, dgamma, dbeta, dy)


def fused_adam(p, g, m, v, step, lr, beta1, beta2, eps, weight_decay, grad_scale=1.0,
               p_lowp=None, zero_grad=False):
    if _native(p):
        bc1 = 1.0 - beta1 ** step
        bc2 = 1.0 - beta2 ** step
        _ops().fused_adam(p, g, m, v, float(lr), float(beta1), float(beta2), float(eps),
                          float(weight_decay), float(grad_scale), float(bc1), float(bc2),
                          p_lowp, bool(zero_grad))
        _tick()
        return
    ref.fused_adam(p, g, m, v, step, lr, beta1, beta2, eps, weight_decay, grad_scale,
                   None if p_lowp is p else p_lowp, zero_grad)


def fused_sgd(p, g, buf, lr, momentum, weight_decay, grad_scale=1.0, p_lowp=None,
              zero_grad=False, first_step=False):
    if _native(p):
        _ops().fused_sgd(p, g, buf, float(lr), float(momentum), float(weight_decay),
                         float(grad_scale), p_lowp, bool(zero_grad), bool(first_step))
        _tick()
        return
    ref.fused_sgd(p, g, buf, lr, momentum, weight_decay, grad_scale,
                  None if p_lowp is p else p_lowp, zero_grad, first_step)


def prep_input(img_u8, out):
    if _native(out):
        _ops().prep_input(img_u8, out)
        _tick()
        return
    ref.prep_input(img_u8, out)


# -- codecs (parallel.codecs) ------------------------------------------------------------------
def codec_amax(flat):
    out = torch.zeros((), dtype=torch.float32, device=flat.device)
    _ops().amax(flat, out)
    _tick()
    return out


def codec_encode(flat, scale, codec):
    payload = torch.empty(flat.numel(), dtype=codec.wire_dtype, device=flat.device)
    _ops().codec_encode(flat, scale, payload.view(torch.uint8), codec.name)
    _tick()
    return payload


def codec_decode(payload, scale, codec, out=None, accumulate=False):
    if out is None:
        out = torch.empty(payload.numel(), dtype=torch.float32, device=payload.device)
        accumulate = False
    _ops().codec_decode(payload.view(torch.uint8), scale, out, codec.name, bool(accumulate))
    _tick()
    return out
