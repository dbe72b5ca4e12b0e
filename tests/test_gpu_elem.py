"""Memory-bound sm_100a kernels against the plain-PyTorch fp32 oracles (ops.reference)."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn, reference as ref, loader
from distributed_deep_learning_on_personal_computers_b200.parallel.codecs import get_codec, reference_restatement

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bf(x):
    return x.to(torch.bfloat16)


def test_native_library_is_loaded():
    assert loader.library() is not None, loader.load_error()


@pytest.mark.parametrize("C,pool", [(64, True), (32, False), (128, True), (16, False), (96, True), (192, False),
                                    (64, False)])
def test_bn_relu_pool_fwd_and_stats(C, pool):
    torch.manual_seed(0)
    N, H, W = (3, 16, 24) if C != 64 else (5, 96, 136)      # the large one runs several trips per thread
    y = _bf(torch.randn(N, H, W, C, device=DEV) * 2 + 0.5)
    part = torch.empty(torch.ops.b200seg.stats_rows(), 2, C, device=DEV)
    torch.ops.b200seg.channel_stats(y, part)
    want = torch.stack([y.float().sum((0, 1, 2)), (y.float() ** 2).sum((0, 1, 2))])
    torch.testing.assert_close(part.sum(0), want, rtol=1e-4, atol=1e-2)
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    nbt = torch.zeros((), dtype=torch.long, device=DEV)
    ss = torch.empty(4, C, device=DEV)
    Fn.bn_finalize(part, N * H * W, gamma, beta, rm, rv, nbt, ss)
    rm2, rv2 = torch.zeros(C), torch.ones(C)
    ss_ref = ref.bn_finalize(want.cpu(), N * H * W, gamma.cpu(), beta.cpu(), rm2, rv2)
    torch.testing.assert_close(ss.cpu(), ss_ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rm.cpu(), rm2, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(rv.cpu(), rv2, rtol=1e-4, atol=1e-6)
    assert nbt.item() == 1
    # outputs go into a channel slice of a wider buffer (the concat trick)
    wide = torch.zeros(N, H, W, C + 32, dtype=torch.bfloat16, device=DEV)
    out = wide[..., 32:]
    pl = torch.empty(N, H // 2, W // 2, C, dtype=torch.bfloat16, device=DEV) if pool else None
    Fn.bn_relu_pool_fwd(y, ss, out, pl)
    o_ref = torch.empty(N, H, W, C, dtype=torch.bfloat16)
    p_ref = torch.empty(N, H // 2, W // 2, C, dtype=torch.bfloat16) if pool else None
    ref.bn_relu_pool_fwd(y.cpu(), ss.cpu(), o_ref, p_ref)
    torch.testing.assert_close(out.cpu().float(), o_ref.float(), rtol=2e-2, atol=2e-2)
    assert (wide[..., :32] == 0).all()
    if pool:
        torch.testing.assert_close(pl.cpu().float(), p_ref.float(), rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("C,pool,with_dout", [(64, True, True), (32, False, True), (128, True, False),
                                              (96, False, True), (192, True, True), (64, False, True)])
def test_bn_relu_pool_bwd(C, pool, with_dout):
    torch.manual_seed(1)
    N, H, W = (2, 16, 16) if C != 64 else (5, 96, 136)
    y = _bf(torch.randn(N, H, W, C, device=DEV))
    gamma = torch.rand(C, device=DEV) + 0.5
    beta = torch.randn(C, device=DEV) * 0.3
    stats = torch.stack([y.float().sum((0, 1, 2)), (y.float() ** 2).sum((0, 1, 2))])
    ss = ref.bn_finalize(stats, N * H * W, gamma, beta, None, None).to(DEV)
    d_out = _bf(torch.randn(N, H, W, C, device=DEV)) if with_dout else None
    d_pool = _bf(torch.randn(N, H // 2, W // 2, C, device=DEV)) if pool else None
    dg, db = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    dy = torch.empty_like(y)
    Fn.bn_relu_pool_bwd(d_out, d_pool, y, ss, gamma, dg, db, None, dy)
    dg_r, db_r = torch.zeros(C), torch.zeros(C)
    dy_r = torch.empty(N, H, W, C)
    ref.bn_relu_pool_bwd(None if d_out is None else d_out.cpu().float(),
                         None if d_pool is None else d_pool.cpu().float(), y.cpu(), ss.cpu(), gamma.cpu(),
                         dg_r, db_r, None, dy_r)
    torch.testing.assert_close(db.cpu(), db_r, rtol=1e-3, atol=5e-2)
    torch.testing.assert_close(dg.cpu(), dg_r, rtol=1e-3, atol=5e-2)
    torch.testing.assert_close(dy.cpu().float(), dy_r, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("C,K", [(64, 6), (32, 6), (128, 3), (64, 8), (128, 7), (48, 5), (256, 12)])
def test_head_ce_fwd_bwd(C, K):
    torch.manual_seed(2)
    N, H, W = 2, 24, 20
    a = _bf(torch.randn(N, H, W, C, device=DEV))
    w = torch.randn(K, C, device=DEV) * 0.2
    b = torch.randn(K, device=DEV) * 0.1
    lab = torch.randint(0, K, (N, H, W), dtype=torch.uint8, device=DEV)
    for rep in range(2):                                  # second call checks the counter reset
        da = torch.empty_like(a)
        dw, db = torch.zeros(K, C, device=DEV), torch.zeros(K, device=DEV)
        pred = torch.empty(N, H, W, dtype=torch.uint8, device=DEV)
        acc = torch.zeros(2, device=DEV)
        scale = 1.0 / (N * H * W)
        Fn.head_ce_fwd_bwd(a, w, b, lab, scale, da, dw, db, pred, acc)
        da_r = torch.empty(N, H, W, C)
        dw_r, db_r = torch.zeros(K, C), torch.zeros(K)
        pred_r = torch.empty(N, H, W, dtype=torch.uint8)
        acc_r = ref.head_ce_fwd_bwd(a.cpu(), w.cpu(), b.cpu(), lab.cpu(), scale, da_r, dw_r, db_r, pred_r)
        torch.testing.assert_close(acc.cpu()[0], acc_r[0], rtol=1e-4, atol=1e-2)
        assert abs(acc.cpu()[1].item() - acc_r[1].item()) <= 2
        assert (pred.cpu() != pred_r).float().mean().item() < 1e-3
        torch.testing.assert_close(da.cpu().float(), da_r, rtol=2e-2, atol=1e-6)
        torch.testing.assert_close(dw.cpu(), dw_r, rtol=1e-3, atol=1e-5)
        torch.testing.assert_close(db.cpu(), db_r, rtol=1e-3, atol=1e-6)
    acc2 = torch.zeros(2, device=DEV)
    Fn.head_ce_fwd_bwd(a, w, b, lab, 0.0, None, None, None, None, acc2)      # eval form
    torch.testing.assert_close(acc2.cpu()[0], acc_r[0], rtol=1e-4, atol=1e-2)


@pytest.mark.parametrize("C,K,shape", [(64, 6, (2, 24, 20)), (32, 6, (1, 7, 9)), (128, 3, (2, 16, 16)),
                                       (64, 8, (1, 40, 33)), (128, 7, (1, 9, 11)), (64, 1, (1, 8, 8))])
def test_head_tail_one_pass(C, K, shape):
    """BN+ReLU on load + head + CE + head backward + BN-backward sums (csrc/elem/head_tail.cu) against the plain
    PyTorch fp32 restatement of the same chain (ops.reference.head_tail_fwd_bwd)."""
    torch.manual_seed(5)
    N, H, W = shape
    P = N * H * W
    assert Fn.head_tail_supported(torch.empty(1, 1, 1, C, device=DEV, dtype=torch.bfloat16), K)
    y = _bf(torch.randn(N, H, W, C, device=DEV) * 1.5)
    ss = torch.stack([torch.rand(C) + 0.5, torch.randn(C) * 0.3, torch.randn(C) * 0.1, torch.rand(C) + 0.5]).to(DEV)
    w = torch.randn(K, C, device=DEV) * 0.2
    b = torch.randn(K, device=DEV) * 0.1
    lab = torch.randint(0, K, (N, H, W), dtype=torch.uint8, device=DEV)
    rows = Fn.head_tail_rows(y)
    scale = 1.0 / P
    g_r = torch.empty(N, H, W, C)
    dw_r, db_r = torch.zeros(K, C), torch.zeros(K)
    pred_r = torch.empty(N, H, W, dtype=torch.uint8)
    part_r = torch.zeros(1, 2, C)
    acc_r = ref.head_tail_fwd_bwd(y.cpu(), ss.cpu(), w.cpu(), b.cpu(), lab.cpu(), scale, g_r, dw_r, db_r, pred_r, part_r)
    for rep in range(2):                                  # second call: counter reset, accumulation into dw / db
        g = torch.full_like(y, float("nan"))
        dw, db = torch.zeros(K, C, device=DEV), torch.zeros(K, device=DEV)
        pred = torch.full((N, H, W), 255, dtype=torch.uint8, device=DEV)
        acc = torch.zeros(2, device=DEV)
        part = torch.full((rows, 2, C), float("nan"), device=DEV)
        Fn.head_tail_fwd_bwd(y, ss, w, b, lab, scale, g, dw, db, pred, acc, part)
        torch.cuda.synchronize()
        torch.testing.assert_close(acc.cpu()[0], acc_r[0], rtol=1e-4, atol=1e-2)
        assert abs(acc.cpu()[1].item() - acc_r[1].item()) <= 2
        assert (pred.cpu() != pred_r).float().mean().item() < 2e-3
        gc = g.cpu().float()
        assert torch.isfinite(gc).all()
        assert (gc - g_r).abs().max().item() <= 1.5e-2 * g_r.abs().max().item() + 1e-9
        torch.testing.assert_close(dw.cpu(), dw_r, rtol=2e-3, atol=2e-3 * dw_r.abs().max().item())
        torch.testing.assert_close(db.cpu(), db_r, rtol=1e-3, atol=1e-6)
        # the BatchNorm-backward sums must be those of the g the kernel STORED (bf16) under the mask of y
        pre = y.cpu().float() * ss.cpu()[0] + ss.cpu()[1]
        gm = gc * (pre > 0)
        want = torch.stack([gm.sum(dim=(0, 1, 2)), (gm * y.cpu().float()).sum(dim=(0, 1, 2))])
        got = part.cpu().double().sum(dim=0).float()
        assert torch.isfinite(got).all()
        tol = 1e-3 * torch.stack([gm.abs().sum(dim=(0, 1, 2)), (gm * y.cpu().float()).abs().sum(dim=(0, 1, 2))]) + 1e-9
        assert ((got - want).abs() <= tol).all()
    acc2 = torch.zeros(2, device=DEV)
    pred2 = torch.empty(N, H, W, dtype=torch.uint8, device=DEV)
    Fn.head_tail_fwd_bwd(y, ss, w, b, lab, 0.0, None, None, None, pred2, acc2, None)      # forward-only form
    torch.testing.assert_close(acc2.cpu()[0], acc_r[0], rtol=1e-4, atol=1e-2)
    assert torch.equal(pred2, pred)
    # finalize + apply from these partial sums == the three-kernel BN backward on the same g
    gamma = (torch.rand(C, device=DEV) + 0.5)
    dy_a, dy_b = torch.empty_like(y), torch.empty_like(y)
    dg_a, dbt_a = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    dg_b, dbt_b = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    Fn.bn_relu_bwd_from_partial(part, rows, g, y, ss, gamma, dg_a, dbt_a, dy_a)
    Fn.bn_relu_pool_bwd(g, None, y, ss, gamma, dg_b, dbt_b, None, dy_b)
    torch.testing.assert_close(dg_a, dg_b, rtol=1e-3, atol=1e-4 * dg_b.abs().max().item() + 1e-7)
    torch.testing.assert_close(dbt_a, dbt_b, rtol=1e-3, atol=1e-4 * dbt_b.abs().max().item() + 1e-7)
    assert (dy_a.float() - dy_b.float()).abs().max().item() <= 1e-2 * dy_b.float().abs().max().item() + 1e-9


def test_fused_adam_matches_torch_adam():
    torch.manual_seed(3)
    n = 64 * 1000
    p0 = torch.randn(n)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=1e-3)
    p, m, v = p0.clone().to(DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    lp = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    for step in range(1, 4):
        g = torch.randn(n) * 0.1
        ref_p.grad = (g * 0.5).clone()
        opt.step()
        gd = g.to(DEV)
        Fn.fused_adam(p, gd, m, v, step, 1e-3, 0.9, 0.999, 1e-8, 0.0, grad_scale=0.5, p_lowp=lp, zero_grad=True)
        assert gd.abs().max().item() == 0.0
    torch.testing.assert_close(p.cpu(), ref_p.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lp.float().cpu(), p.cpu().to(torch.bfloat16).float())


def test_fused_sgd_momentum():
    torch.manual_seed(4)
    n = 64 * 100
    p0 = torch.randn(n)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=1e-3)
    p, buf = p0.clone().to(DEV), torch.zeros(n, device=DEV)
    for step in range(1, 4):
        g = torch.randn(n)
        ref_p.grad = g.clone()
        opt.step()
        Fn.fused_sgd(p, g.to(DEV), buf, 0.1, 0.9, 1e-3, first_step=(step == 1))
    torch.testing.assert_close(p.cpu(), ref_p.detach(), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["fp32", "bf16", "fp16", "q201", "q21", "int8s", "fp8"])
def test_codecs_match_cpu_oracle(name):
    torch.manual_seed(5)
    g = (torch.randn(100_003) * 0.01).to(DEV)
    c = get_codec(name)
    q, s = c.encode(g)
    qc, sc = c.encode(g.cpu())
    assert s.item() == sc.item()
    if name == "fp8":
        assert (q.float().cpu() - qc.float()).abs().max() <= 32        # one fp8 ulp at the top
    else:
        assert torch.equal(q.cpu(), qc)
    back = c.decode(q, s)
    torch.testing.assert_close(back.cpu(), c.decode(qc, sc), rtol=0, atol=0 if name != "fp8" else 1e-3)
    if name in ("q201", "q21"):
        assert torch.equal(back.cpu(), reference_restatement(g.cpu(), "float16" if name == "q201" else "int8"))
    acc = torch.ones_like(g)
    c.decode(q, s, out=acc, accumulate=True)
    torch.testing.assert_close(acc, back + 1)


def test_codec_amax_kernel():
    g = torch.randn(12345, device=DEV)
    assert Fn.codec_amax(g).item() == g.abs().max().item()
    assert Fn.codec_amax(torch.zeros(100, device=DEV)).item() == 0.0


def test_prep_input():
    img = torch.randint(0, 256, (2, 8, 8, 3), dtype=torch.uint8, device=DEV)
    out = torch.empty(2, 8, 8, 16, dtype=torch.bfloat16, device=DEV)
    Fn.prep_input(img, out)
    want = torch.zeros(2, 8, 8, 16)
    ref.prep_input(img.cpu(), want)
    torch.testing.assert_close(out.cpu().float(), want.to(torch.bfloat16).float())


def test_bilinear_upsample_fwd_bwd():
    torch.manual_seed(6)
    x = _bf(torch.randn(2, 5, 7, 16, device=DEV))
    wide = torch.zeros(2, 10, 14, 24, dtype=torch.bfloat16, device=DEV)
    Fn.upsample_bilinear2x_fwd(x, wide[..., :16])
    want = torch.empty(2, 10, 14, 16)
    ref.upsample_bilinear2x_fwd(x.cpu().float(), want)
    torch.testing.assert_close(wide[..., :16].cpu().float(), want, rtol=2e-2, atol=2e-2)
    g = _bf(torch.randn(2, 10, 14, 16, device=DEV))
    dx = torch.empty_like(x)
    Fn.upsample_bilinear2x_bwd(g, dx)
    dx_r = torch.empty(2, 5, 7, 16)
    ref.upsample_bilinear2x_bwd(g.cpu().float(), dx_r)
    torch.testing.assert_close(dx.cpu().float(), dx_r, rtol=2e-2, atol=3e-2)
