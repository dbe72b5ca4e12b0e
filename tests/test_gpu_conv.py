"""tcgen05 implicit-GEMM convolutions against PyTorch fp32 (the reference's own math)."""
import pytest
import torch
import torch.nn.functional as F

from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn, reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"

SHAPES = [  # N, H, W, Cin, Cout
    (2, 32, 32, 64, 64),
    (1, 64, 48, 64, 128),
    (2, 16, 16, 128, 256),
    (4, 8, 8, 256, 512),
    (1, 32, 32, 192, 64),
    (2, 32, 32, 16, 64),       # stem: 3 channels padded to 16
    (1, 32, 32, 32, 32),       # reference width 32
    (1, 16, 16, 96, 32),
    (3, 12, 20, 64, 64),       # ragged tiles
    (1, 6, 256, 64, 64),       # strip mode (rows of 128-pixel segments), resident filter bank
    (2, 5, 128, 192, 64),      # strip mode, streamed weights, 3 channel chunks
    (1, 4, 128, 128, 128),     # strip mode, N = 128
    (1, 8, 128, 16, 64),       # strip mode, stem (K chunk of 16)
    (2, 6, 256, 128, 64),      # strip mode, two-row work units (streamed weights, even height)
    (2, 12, 256, 64, 64),      # strip mode, resident bank, four-row units with N-merged vertical taps
    (1, 4, 128, 64, 64),       # the same, a single unit
    (1, 40, 64, 192, 64),      # wgrad strip ring wraps (mirror slot), two pieces per column
    (2, 40, 128, 64, 256),     # wgrad strip, 128-column variant + tap (2,2) on the split-K kernel
    (2, 8, 128, 192, 64),      # dgrad: three 64-column N tiles, each with its resident bank slice and four-row units
    (2, 8, 256, 64, 128),      # fprop: two such N tiles (64 -> 128), register statistics per tile
    (1, 8, 256, 128, 256),     # strip mode, 256 output channels as two 128-column N tiles with two-row units
]


def _mk(N, H, W, Cin, Cout, seed=0):
    torch.manual_seed(seed)
    x = (torch.randn(N, H, W, Cin, device=DEV)).to(torch.bfloat16)
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV) * (1.0 / (3 * Cin ** 0.5))).to(torch.bfloat16)
    b = torch.randn(Cout, device=DEV)
    return x, w, b


@pytest.mark.parametrize("env", [{"B200SEG_ROWS4": "0"}, {"B200SEG_EPI_PRIVATE": "0"}, {"B200SEG_WGRAD_PAIR": "0"}])
def test_fprop_alternative_paths(env):
    """The defaults are four-row N-merged units, the warp-private epilogue and the row-pair wgrad; the paths
    they replaced stay selectable (A/B runs) and must stay correct.  The switches are read once per process,
    hence a subprocess."""
    import subprocess
    import sys
    code = (
        "import torch\n"
        "from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn, reference as ref\n"
        "torch.manual_seed(0)\n"
        "for (N,H,W,Ci,Co) in ((2,12,256,64,64),(1,8,128,16,64)):\n"
        "    x=torch.randn(N,H,W,Ci,device='cuda').bfloat16(); w=(torch.randn(Co,3,3,Ci,device='cuda')/(3*Ci**0.5)).bfloat16()\n"
        "    b=torch.randn(Co,device='cuda'); y=torch.empty(N,H,W,Co,dtype=torch.bfloat16,device='cuda')\n"
        "    part=Fn.conv3x3_fwd(x,w,b,y,want_stats=True,impl='tcgen05')\n"
        "    yr,_=ref.conv3x3_fwd(x.float(),w.float(),b,want_stats=False)\n"
        "    torch.testing.assert_close(y.float(),yr,rtol=2e-2,atol=2e-2)\n"
        "    torch.testing.assert_close(part.sum(0)[0],y.float().sum((0,1,2)),rtol=2e-3,atol=0.5)\n"
        "    dy=(torch.randn(N,H,W,Co,device='cuda')*0.1).bfloat16(); dw=torch.zeros(Co,3,3,Ci,device='cuda'); dr=torch.zeros_like(dw)\n"
        "    Fn.conv3x3_wgrad(x,dy,dw,impl='tcgen05'); ref.conv3x3_wgrad(x.float(),dy.float(),dr)\n"
        "    assert (dw-dr).abs().max().item() <= 1e-2*dr.abs().max().item()+1e-3\n"
        "print('ok')\n")
    import os
    env = dict(os.environ, **env)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


@pytest.mark.parametrize("shape", SHAPES)
def test_conv3x3_fprop_and_stats(shape):
    N, H, W, Cin, Cout = shape
    x, w, b = _mk(*shape)
    wide = torch.zeros(N, H, W, Cin + 16, dtype=torch.bfloat16, device=DEV)       # input as a channel slice
    wide[..., :Cin] = x
    y = torch.empty(N, H, W, Cout, dtype=torch.bfloat16, device=DEV)
    part = Fn.conv3x3_fwd(wide[..., :Cin], w, b, y, want_stats=True, impl="tcgen05")
    y_ref, _ = ref.conv3x3_fwd(x.float(), w.float(), b, want_stats=False)
    torch.testing.assert_close(y.float(), y_ref, rtol=2e-2, atol=2e-2)
    st = part.sum(0)
    # the statistics are those of the conv output: of its bf16 rounding for wide tiles, of the fp32
    # accumulators for 64-channel tiles (register path); both are within rounding noise of each other
    want = torch.stack([y.float().sum((0, 1, 2)), (y.float() ** 2).sum((0, 1, 2))])
    noise = 0.02 * (N * H * W) ** 0.5
    torch.testing.assert_close(st, want, rtol=2e-3, atol=noise)
    want32 = torch.stack([y_ref.sum((0, 1, 2)), (y_ref ** 2).sum((0, 1, 2))])
    torch.testing.assert_close(st, want32, rtol=2e-3, atol=noise)


@pytest.mark.parametrize("shape", [s for s in SHAPES if s[3] != 16])
def test_conv3x3_dgrad(shape):
    N, H, W, Cin, Cout = shape
    _, w, _ = _mk(*shape)
    dy = torch.randn(N, H, W, Cout, device=DEV).to(torch.bfloat16)
    wide = torch.zeros(N, H, W, Cin + 32, dtype=torch.bfloat16, device=DEV)
    dx = wide[..., 32:]
    Fn.conv3x3_dgrad(dy, w, dx, impl="tcgen05")
    dx_ref = ref.conv3x3_dgrad(dy.float(), w.float())
    torch.testing.assert_close(dx.float(), dx_ref, rtol=2e-2, atol=3e-2)
    assert (wide[..., :32] == 0).all()
    # same kernel with column sums of dx from the epilogue (the transposed conv's bias gradient)
    dx2 = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device=DEV)
    part = Fn.conv3x3_dgrad_stats(dy, w, dx2, impl="tcgen05")
    assert torch.equal(dx2, dx.contiguous())
    want = torch.stack([dx2.float().sum((0, 1, 2)), (dx2.float() ** 2).sum((0, 1, 2))])
    # 64-column tiles sum the fp32 accumulators, wider ones the bf16-rounded tile: allow rounding noise
    # (2^-9 per element, random sign) on top of the summation order.  The column sums are what the engine uses
    # (the transposed conv's bias gradient); the warp-private epilogue of the strip layers leaves the
    # sum-of-squares row at zero for a dgrad.
    got = part.sum(0)
    torch.testing.assert_close(got[0], want[0], rtol=2e-3, atol=0.02 * (N * H * W) ** 0.5)
    if float(got[1].abs().sum()) > 0:
        torch.testing.assert_close(got[1], want[1], rtol=2e-3, atol=0.02 * (N * H * W) ** 0.5)


@pytest.mark.parametrize("shape", SHAPES)
def test_conv3x3_wgrad(shape):
    N, H, W, Cin, Cout = shape
    x, w, _ = _mk(*shape)
    dy = (torch.randn(N, H, W, Cout, device=DEV) * 0.1).to(torch.bfloat16)
    dw = torch.ones(Cout, 3, 3, Cin, device=DEV)              # accumulates on top of existing values
    Fn.conv3x3_wgrad(x, dy, dw, impl="tcgen05")
    dw_ref = torch.ones(Cout, 3, 3, Cin, device=DEV)
    ref.conv3x3_wgrad(x.float(), dy.float(), dw_ref)
    scale = (dw_ref - 1).abs().max().item()
    assert (dw - dw_ref).abs().max().item() <= 1e-2 * scale + 1e-3


@pytest.mark.parametrize("N,h,w,C", [(2, 16, 16, 64), (1, 8, 8, 256), (2, 32, 32, 128), (4, 4, 4, 64)])
def test_convT2x2_fwd_into_concat_slice(N, h, w, C):
    torch.manual_seed(7)
    x = torch.randn(N, h, w, C, device=DEV).to(torch.bfloat16)
    wp = (torch.randn(C, C, 2, 2, device=DEV) * 0.05)                    # logical (Cin, Cout, 2, 2)
    w_lp = wp.permute(2, 3, 1, 0).contiguous().view(4 * C, C).to(torch.bfloat16)
    bias = torch.randn(C, device=DEV)
    cat = torch.zeros(N, 2 * h, 2 * w, C + 64, dtype=torch.bfloat16, device=DEV)
    Fn.convT2x2_fwd(x, wp, w_lp, bias, cat[..., :C], impl="tcgen05")
    want = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), w_lp.float().view(2, 2, C, C).permute(3, 2, 0, 1),
                              bias, stride=2).permute(0, 2, 3, 1)
    torch.testing.assert_close(cat[..., :C].float(), want, rtol=2e-2, atol=2e-2)
    assert (cat[..., C:] == 0).all()


@pytest.mark.parametrize("N,h,w,C", [(2, 16, 16, 64), (1, 8, 8, 256), (2, 32, 32, 128), (4, 4, 4, 64)])
def test_convT2x2_bwd(N, h, w, C):
    torch.manual_seed(8)
    x = torch.randn(N, h, w, C, device=DEV).to(torch.bfloat16)
    wp = torch.randn(C, C, 2, 2, device=DEV) * 0.05
    w_lp = wp.permute(2, 3, 1, 0).contiguous().view(4 * C, C).to(torch.bfloat16)
    dcat = (torch.randn(N, 2 * h, 2 * w, C + 64, device=DEV) * 0.1).to(torch.bfloat16)
    d_up = dcat[..., :C]                                   # gradient arrives as a concat-buffer slice
    flat = torch.ones(4 * C * C, device=DEV)
    dw_param = flat.view(2, 2, C, C).permute(3, 2, 0, 1)     # logical (Cin, Cout, 2, 2) view
    dbias = torch.zeros(C, device=DEV)
    dx = torch.empty(N, h, w, C, dtype=torch.bfloat16, device=DEV)
    Fn.convT2x2_bwd(x, d_up, wp, w_lp, dw_param, dbias, dx, impl="tcgen05")
    wl = w_lp.float().view(2, 2, C, C).permute(3, 2, 0, 1)
    dw_ref = torch.ones(C, C, 2, 2, device=DEV)
    db_ref = torch.zeros(C, device=DEV)
    dx_ref = torch.empty(N, h, w, C, device=DEV)
    ref.convT2x2_bwd(x.float(), d_up.float(), wl, dw_ref, db_ref, dx_ref)
    torch.testing.assert_close(dx.float(), dx_ref, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(dbias, db_ref, rtol=1e-3, atol=1e-2)
    scale = (dw_ref - 1).abs().max().item()
    assert (dw_param - dw_ref).abs().max().item() <= 1e-2 * scale + 1e-3


# ---- the shapes that bench.py times (headline config: U-Net b=64, batch 16, 512x512) ----------------------
# Several waves of tiles per SM, double-buffered accumulators wrapping many times, full-size rings.
BENCH_SHAPES = [  # N, H, W, Cin, Cout
    (16, 512, 512, 64, 64),      # L0.c1 / L0.dec1: resident bank, four-row units, warp-private epilogue
    (16, 512, 512, 192, 64),     # L0.dec0: streamed weights, two-row units
    (16, 512, 512, 16, 64),      # stem
    (16, 256, 256, 384, 128),    # L1.dec0
    (16, 32, 32, 1024, 512),     # L4.dec0: box mode, split-K wgrad in CTA pairs
    (16, 16, 16, 512, 512),      # L5
]


def _close_big(got, want, rtol, atol):
    """assert_close without materialising several copies of a multi-GB tensor."""
    err = (got.float() - want).abs_()
    tol = want.abs().mul_(rtol).add_(atol)
    bad = (err > tol).sum().item()
    assert bad == 0, f"{bad} elements out of tolerance, max abs err {err.max().item():.4g}"


@pytest.mark.parametrize("shape", BENCH_SHAPES)
def test_bench_shapes_fprop_dgrad_wgrad(shape):
    N, H, W, Cin, Cout = shape
    x, w, _ = _mk(*shape, seed=3)
    xn = x.float().permute(0, 3, 1, 2)
    wn = w.float().permute(0, 3, 1, 2)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False                      # the oracle is plain fp32
    try:
        # fprop + BN statistics
        y = torch.empty(N, H, W, Cout, dtype=torch.bfloat16, device=DEV)
        part = Fn.conv3x3_fwd(x, w, None, y, want_stats=True, impl="tcgen05")
        y_ref = F.conv2d(xn, wn, None, padding=1).permute(0, 2, 3, 1)
        _close_big(y, y_ref, 2e-2, 2e-2)
        want = torch.stack([y_ref.sum((0, 1, 2)), (y_ref ** 2).sum((0, 1, 2))])
        torch.testing.assert_close(part.sum(0), want, rtol=3e-3, atol=0.03 * (N * H * W) ** 0.5)
        del y_ref
        dy = (torch.randn(N, H, W, Cout, device=DEV) * 0.1).to(torch.bfloat16)
        dyn = dy.float().permute(0, 3, 1, 2)
        # wgrad (accumulating)
        dw = torch.ones(Cout, 3, 3, Cin, device=DEV)
        Fn.conv3x3_wgrad(x, dy, dw, impl="tcgen05")
        dw_ref = torch.nn.grad.conv2d_weight(xn, wn.shape, dyn, padding=1).permute(0, 2, 3, 1)
        scale = dw_ref.abs().max().item()
        assert (dw - 1 - dw_ref).abs().max().item() <= 1e-2 * scale + 1e-3
        # dgrad
        if Cin != 16:
            dx = torch.empty(N, H, W, Cin, dtype=torch.bfloat16, device=DEV)
            Fn.conv3x3_dgrad(dy, w, dx, impl="tcgen05")
            dx_ref = torch.nn.grad.conv2d_input(xn.shape, wn, dyn, padding=1).permute(0, 2, 3, 1)
            _close_big(dx, dx_ref, 2e-2, 3e-2)
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_bench_shape_convT():
    N, h, w, C = 16, 256, 256, 128                                # up_conv1.up_sample of the headline config
    torch.manual_seed(9)
    x = torch.randn(N, h, w, C, device=DEV).to(torch.bfloat16)
    wp = torch.randn(C, C, 2, 2, device=DEV) * 0.05
    w_lp = wp.permute(2, 3, 1, 0).contiguous().view(4 * C, C).to(torch.bfloat16)
    bias = torch.randn(C, device=DEV)
    cat = torch.zeros(N, 2 * h, 2 * w, C + 64, dtype=torch.bfloat16, device=DEV)
    Fn.convT2x2_fwd(x, wp, w_lp, bias, cat[..., :C], impl="tcgen05")
    want = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), w_lp.float().view(2, 2, C, C).permute(3, 2, 0, 1),
                              bias, stride=2).permute(0, 2, 3, 1)
    _close_big(cat[..., :C], want, 2e-2, 2e-2)
    assert (cat[..., C:] == 0).all()
