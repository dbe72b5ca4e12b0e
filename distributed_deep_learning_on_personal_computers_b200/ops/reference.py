"""Plain-PyTorch statements of every fused op (the oracles, and the CPU path).

Each function has the same signature and the same NHWC buffer conventions as its sm_100a kernel
in ``csrc/``; ``ops.functional`` picks one or the other.  The math is the reference's own: its
device ops are stock PyTorch modules, so "PyTorch fp32 on CPU" *is* the
reference's code path.  Citations are to ``R`` =
``/root/reference/Vaihingen PyTorch 2 (кластер).py``.

Conventions
-----------
* activations: ``(N, H, W, C)`` tensors, possibly channel-slices of a wider buffer (so the last
  dim is contiguous but the pixel stride may exceed ``C``) -- this is how ``torch.cat`` (R:616)
  is eliminated: producers write straight into their slice of the concat buffer;
* conv weights: ``(Cout, 3, 3, Cin)`` = the flat-buffer memory order (``models.unet``);
* statistics and BN vectors are fp32.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def _nchw(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 3, 1, 2)


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 2, 3, 1)


# ---------------------------------------------------------------------------------------------
# K1 / K5(stats): 3x3 conv, pad 1, + bias; also returns per-channel sum and sum of squares of
# the fp32 result (what the kernel's epilogue accumulates from its TMEM accumulators).  R:579.
def conv3x3_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor],
                want_stats: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    cin = w.shape[3]
    wt = w.permute(0, 3, 1, 2)
    xin = _nchw(x[..., :cin])
    y32 = F.conv2d(xin.float(), wt.float(), None if bias is None else bias.float(), padding=1)
    y = _nhwc(y32).contiguous()
    stats = None
    if want_stats:
        stats = torch.stack([y.sum(dim=(0, 1, 2), dtype=torch.float32),
                             (y * y).sum(dim=(0, 1, 2), dtype=torch.float32)])
    return y.to(x.dtype), stats


# K2: data gradient of the 3x3 conv.
def conv3x3_dgrad(dy: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None
                  ) -> torch.Tensor:
    wt = w.permute(0, 3, 1, 2).float()
    dx = F.conv_transpose2d(_nchw(dy).float(), wt, padding=1)
    dx = _nhwc(dx)
    if out is None:
        return dx.contiguous().to(dy.dtype)
    out.copy_(dx)
    return out


# K3: weight gradient, accumulated (+=) in fp32 into ``dw`` of shape (Cout, 3, 3, Cin).
def conv3x3_wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, accumulate: bool = True
                  ) -> torch.Tensor:
    cout, _, _, cin = dw.shape
    xin = _nchw(x[..., :cin]).float()
    g = torch.nn.grad.conv2d_weight(xin, (cout, cin, 3, 3), _nchw(dy).float(), padding=1)
    g = g.permute(0, 2, 3, 1)
    if accumulate:
        dw.add_(g)
    else:
        dw.copy_(g)
    return dw


# K5: finish the batch statistics and fold BN into one scale/shift pair; update running stats
# with momentum 0.1 and the unbiased variance, as nn.BatchNorm2d does (R:580).
def bn_finalize(stats: torch.Tensor, count: int, gamma: torch.Tensor, beta: torch.Tensor,
                running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                momentum: float = 0.1, eps: float = 1e-5, training: bool = True,
                conv_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Returns fp32 (4, C): scale, shift, mean, invstd.

    ``conv_bias`` is the bias of the conv that produced ``y`` when that conv did NOT add it (the
    engine's convention): in training mode a per-channel constant cancels inside the
    normalisation and only shifts the running mean; in eval mode it is folded into ``shift``."""
    cb = 0.0 if conv_bias is None else conv_bias.to(torch.float64)
    if training:
        s = stats.to(torch.float64)
        mean = s[0] / count
        var = (s[1] / count - mean * mean).clamp_(min=0.0)
        if running_mean is not None:
            unbiased = var * (count / max(count - 1, 1))
            running_mean.mul_(1 - momentum).add_((mean + cb).to(running_mean.dtype), alpha=momentum)
            running_var.mul_(1 - momentum).add_(unbiased.to(running_var.dtype), alpha=momentum)
    else:
        mean = running_mean.to(torch.float64) - cb
        var = running_var.to(torch.float64)
    invstd = torch.rsqrt(var + eps)
    scale = gamma.to(torch.float64) * invstd
    shift = beta.to(torch.float64) - mean * scale
    return torch.stack([scale, shift, mean, invstd]).to(torch.float32)


# K5+K7+K8: y -> relu(y*scale+shift) written into ``out`` (a slice of a wider buffer is fine)
# and, when ``pool`` is given, its 2x2 max-pool (R:580-581, R:595).
def bn_relu_pool_fwd(y: torch.Tensor, ss: torch.Tensor, out: torch.Tensor,
                     pool: Optional[torch.Tensor] = None) -> None:
    a = torch.relu(y.float() * ss[0] + ss[1]).to(out.dtype)
    out.copy_(a)
    if pool is not None:
        n, h, w, c = a.shape
        pool.copy_(a.view(n, h // 2, 2, w // 2, 2, c).amax(dim=(2, 4)))


# K6+K7+K8 backward.  ``d_out`` is the gradient w.r.t. the activated output (may be a slice),
# ``d_pool`` the gradient w.r.t. the pooled output (routed to the FIRST maximum of each window
# in row-major order, as ATen's max_pool2d does), ``y`` the raw conv output of the forward.
def bn_relu_pool_bwd(d_out: Optional[torch.Tensor], d_pool: Optional[torch.Tensor],
                     y: torch.Tensor, ss: torch.Tensor, gamma: torch.Tensor,
                     dgamma: torch.Tensor, dbeta: torch.Tensor, dbias: Optional[torch.Tensor],
                     dy: torch.Tensor) -> None:
    scale, shift, mean, invstd = ss[0], ss[1], ss[2], ss[3]
    yf = y.float()
    pre = yf * scale + shift
    a = torch.relu(pre).to(y.dtype)                     # what the forward stored
    g = torch.zeros_like(yf) if d_out is None else d_out.float().clone()
    if d_pool is not None:
        n, h, w, c = a.shape
        win = a.view(n, h // 2, 2, w // 2, 2, c).permute(0, 1, 3, 5, 2, 4).reshape(n, h // 2, w // 2, c, 4)
        first = win.float().argmax(dim=-1, keepdim=True)          # first max in (dy, dx) order
        onehot = torch.zeros_like(win, dtype=torch.float32).scatter_(-1, first, 1.0)
        up = onehot * d_pool.float().unsqueeze(-1)
        up = up.view(n, h // 2, w // 2, c, 2, 2).permute(0, 1, 4, 2, 5, 3).reshape(n, h, w, c)
        g = g + up
    g = g * (pre > 0)
    m = float(yf.shape[0] * yf.shape[1] * yf.shape[2])
    xhat = (yf - mean) * invstd
    sum_g = g.sum(dim=(0, 1, 2))
    sum_gx = (g * xhat).sum(dim=(0, 1, 2))
    dbeta.add_(sum_g)
    dgamma.add_(sum_gx)
    d = (gamma.float() * invstd) * (g - sum_g / m - xhat * (sum_gx / m))
    d = d.to(dy.dtype)
    dy.copy_(d)
    if dbias is not None:
        dbias.add_(d.float().sum(dim=(0, 1, 2)))


# K9+K10: ConvTranspose2d(k=2, s=2) written into the first channels of the concat buffer (R:607,
# R:616).  ``w`` is the logical (Cin, Cout, 2, 2) parameter (any strides).
def convT2x2_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, out: torch.Tensor) -> None:
    up = F.conv_transpose2d(_nchw(x).float(), w.float(), bias.float(), stride=2)
    out.copy_(_nhwc(up))


def convT2x2_bwd(x: torch.Tensor, d_up: torch.Tensor, w: torch.Tensor,
                 dw: torch.Tensor, dbias: torch.Tensor, dx: Optional[torch.Tensor]) -> None:
    g = _nchw(d_up).float()
    dbias.add_(g.sum(dim=(0, 2, 3)))
    # dW[ci, co, dy, dx] = sum_p x[p, ci] * g[p*2 + (dy,dx), co]
    n, h, wd, ci = x.shape
    co = w.shape[1]
    gq = d_up.float().reshape(n, h, 2, wd, 2, co)
    dw.add_(torch.einsum("nhwi,nhywxo->ioyx", x.float(), gq))
    if dx is not None:
        dx.copy_(torch.einsum("nhywxo,ioyx->nhwi", gq, w.float()))


# K11: bilinear x2 with align_corners=True (R:609), into the concat buffer; and its adjoint.
def upsample_bilinear2x_fwd(x: torch.Tensor, out: torch.Tensor) -> None:
    up = F.interpolate(_nchw(x).float(), scale_factor=2, mode="bilinear", align_corners=True)
    out.copy_(_nhwc(up))


def upsample_bilinear2x_bwd(d_up: torch.Tensor, dx: torch.Tensor) -> None:
    n, h2, w2, c = d_up.shape
    with torch.enable_grad():
        probe = torch.zeros(n, c, h2 // 2, w2 // 2, dtype=torch.float32, device=d_up.device,
                            requires_grad=True)
        up = F.interpolate(probe, scale_factor=2, mode="bilinear", align_corners=True)
        (g,) = torch.autograd.grad(up, probe, _nchw(d_up).float())
    dx.copy_(_nhwc(g))


# K12+K13+K20: 1x1 head, softmax cross-entropy (mean over all pixels of the micro-batch, R:703),
# arg-max / correct count (R:775) and the whole backward of the head in one pass.
def head_ce_fwd_bwd(a: torch.Tensor, w: torch.Tensor, b: torch.Tensor, labels: torch.Tensor,
                    grad_scale: float, da: Optional[torch.Tensor], dw: Optional[torch.Tensor],
                    db: Optional[torch.Tensor], pred: Optional[torch.Tensor] = None
                    ) -> torch.Tensor:
    """Returns fp32 (2,): [sum of per-pixel losses, number of correctly classified pixels].

    ``grad_scale`` multiplies d(loss_sum)/d(logits); pass ``1 / pixels`` for the mean loss.
    ``w`` is (classes, C) fp32.
    """
    af = a.float()
    logits = af @ w.float().t() + b.float()
    lab = labels.long()
    logp = torch.log_softmax(logits, dim=-1)
    loss_sum = -logp.gather(-1, lab.unsqueeze(-1)).sum()
    am = logits.argmax(dim=-1)
    correct = (am == lab).sum()
    if pred is not None:
        pred.copy_(am)
    if da is not None:
        dl = torch.softmax(logits, dim=-1)
        dl.scatter_add_(-1, lab.unsqueeze(-1), -torch.ones_like(dl[..., :1]))
        dl = dl * grad_scale
        da.copy_(dl @ w.float())
        dw.add_(torch.einsum("nhwk,nhwc->kc", dl, af))
        db.add_(dl.sum(dim=(0, 1, 2)))
    return torch.stack([loss_sum.float(), correct.float()])


# The tail in one pass (csrc/elem/head_tail.cu): BN+ReLU of the last stage applied on load, the head with its
# loss and backward, and the two per-channel sums of that BN's backward.  ``bn_partial`` is (rows, 2, C).
def head_tail_fwd_bwd(y: torch.Tensor, ss: torch.Tensor, w: torch.Tensor, b: torch.Tensor, labels: torch.Tensor,
                      grad_scale: float, g_out: Optional[torch.Tensor], dw: Optional[torch.Tensor],
                      db: Optional[torch.Tensor], pred: Optional[torch.Tensor],
                      bn_partial: Optional[torch.Tensor]) -> torch.Tensor:
    pre = y.float() * ss[0] + ss[1]
    a = torch.relu(pre).to(y.dtype)                     # what bn_relu_pool_fwd would have stored
    out = head_ce_fwd_bwd(a, w, b, labels, grad_scale, g_out, dw, db, pred)
    if g_out is not None:
        g = g_out.float() * (pre > 0)
        bn_partial.zero_()
        bn_partial[0, 0] = g.sum(dim=(0, 1, 2))
        bn_partial[0, 1] = (g * y.float()).sum(dim=(0, 1, 2))
    return out


def bn_relu_bwd_from_partial(partial: torch.Tensor, rows: int, d_out: torch.Tensor, y: torch.Tensor,
                             ss: torch.Tensor, gamma: torch.Tensor, dgamma: torch.Tensor, dbeta: torch.Tensor,
                             dy: torch.Tensor) -> None:
    """``bn_relu_pool_bwd`` (no pool) with ``sum g`` and ``sum g*y`` given as per-CTA partial rows."""
    scale, shift, mean, invstd = ss[0], ss[1], ss[2], ss[3]
    yf = y.float()
    g = d_out.float() * ((yf * scale + shift) > 0)
    m = float(yf.shape[0] * yf.shape[1] * yf.shape[2])
    sums = partial[:rows].to(torch.float64).sum(dim=0)
    sum_g = sums[0].float()
    sum_gx = (invstd.double() * (sums[1] - mean.double() * sums[0])).float()
    dbeta.add_(sum_g)
    dgamma.add_(sum_gx)
    xhat = (yf - mean) * invstd
    dy.copy_(((gamma.float() * invstd) * (g - sum_g / m - xhat * (sum_gx / m))).to(dy.dtype))


# K17/K18: Adam / SGD over the flat buffers (torch.optim semantics, R:704, R:437), optionally
# also refreshing the bf16 copy of the weights that the conv kernels read.
def fused_adam(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int,
               lr: float, beta1: float, beta2: float, eps: float, weight_decay: float,
               grad_scale: float = 1.0, p_lowp: Optional[torch.Tensor] = None,
               zero_grad: bool = False) -> None:
    gs = g * grad_scale
    if weight_decay:
        gs = gs + weight_decay * p
    m.mul_(beta1).add_(gs, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(gs, gs, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    if p_lowp is not None:
        p_lowp.copy_(p)
    if zero_grad:
        g.zero_()


def fused_sgd(p: torch.Tensor, g: torch.Tensor, buf: Optional[torch.Tensor], lr: float,
              momentum: float, weight_decay: float, grad_scale: float = 1.0,
              p_lowp: Optional[torch.Tensor] = None, zero_grad: bool = False,
              first_step: bool = False) -> None:
    gs = g * grad_scale
    if weight_decay:
        gs = gs + weight_decay * p
    if momentum and buf is not None:
        if first_step:
            buf.copy_(gs)
        else:
            buf.mul_(momentum).add_(gs)
        gs = buf
    p.add_(gs, alpha=-lr)
    if p_lowp is not None:
        p_lowp.copy_(p)
    if zero_grad:
        g.zero_()


# K19: uint8 HWC tiles -> normalised NHWC activations with the channel dim zero-padded (R:737).
def prep_input(img_u8: torch.Tensor, out: torch.Tensor) -> None:
    c = img_u8.shape[-1]
    out.zero_()
    out[..., :c].copy_(img_u8.float() / 255.0)
