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
]


def _mk(N, H, W, Cin, Cout, seed=0):
    torch.manual_seed(seed)
    x = (torch.randn(N, H, W, Cin, device=DEV)).to(torch.bfloat16)
    w = (torch.randn(Cout, 3, 3, Cin, device=DEV) * (1.0 / (3 * Cin ** 0.5))).to(torch.bfloat16)
    b = torch.randn(Cout, device=DEV)
    return x, w, b


def test_fprop_four_row_units_forced(monkeypatch):
    """The N-merged four-row units are the default for dgrad only; fprop takes them with B200SEG_ROWS4=2,
    which has to be set before the library reads it, hence a subprocess."""
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
        "print('ok')\n")
    import os
    env = dict(os.environ, B200SEG_ROWS4="2")
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
    # (2^-9 per element, random sign) on top of the summation order
    torch.testing.assert_close(part.sum(0), want, rtol=2e-3, atol=0.02 * (N * H * W) ** 0.5)


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
