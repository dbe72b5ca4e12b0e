"""End-to-end on the GPU: native engine vs eager autograd, CUDA-graph step, loss decreases."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig, UNet
from distributed_deep_learning_on_personal_computers_b200.models.engine import NativeEngine
from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("conv,tail", [("cudnn", False), ("tcgen05", False), ("tcgen05", True)])
def test_engine_gradients_match_oracle(conv, tail):
    """GPU engine (sm_100a kernels) vs the SAME plan run by the plain-PyTorch oracles on the CPU with
    bf16 storage between ops (identical rounding points), and vs fp32 eager autograd (sanity)."""
    torch.manual_seed(0)
    own = UNet(out_classes=6, base_channels=32).cuda()
    cpu = UNet(out_classes=6, base_channels=32)
    ref = UNet(out_classes=6, base_channels=32).cuda()
    cpu.load_state_dict(own.state_dict())
    ref.load_state_dict(own.state_dict())
    own.flatten_parameters(torch.device("cuda"))
    cpu.flatten_parameters(torch.device("cpu"))
    N, H, W = 4, 128, 128
    img = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
    lab = torch.randint(0, 6, (N, H, W), dtype=torch.uint8, device="cuda")
    eng = NativeEngine(own, N, H, W, conv_impl=conv, fused_tail=tail)   # tail: head_tail.cu instead of three kernels
    assert eng.fused_tail == tail
    eng.load_batch(img, lab)
    out = eng.train_microbatch().clone()
    eng_cpu = NativeEngine(cpu, N, H, W, act_dtype=torch.bfloat16, fused_tail=False)
    eng_cpu.load_batch(img.cpu(), lab.cpu())
    out_cpu = eng_cpu.train_microbatch().clone()
    assert abs(out[0].item() - out_cpu[0].item()) / out_cpu[0].item() < 2e-2
    worst = 1.0
    table = []
    for (n, p), (_, q) in zip(own.named_parameters(), cpu.named_parameters()):
        if n.endswith((".0.bias", ".3.bias")) and "double_conv" in n:
            continue
        a, b = p.grad.float().flatten().cpu(), q.grad.float().flatten()
        cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
        worst = min(worst, cos)
        table.append(f"{cos:.4f} {n}")
        # gradients near the loss must agree tightly; further back bf16 rounding is amplified by
        # the BatchNorms of the 4x4 / 8x8 levels (few samples per channel), so only the direction
        if n.startswith(("conv_last", "up_conv1.double_conv")):
            assert cos > 0.99, (n, cos)
        assert cos > 0.85, (n, cos)
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    open(f"gpurun_out/grad_cosines_{conv}{'_tail' if tail else ''}.txt", "w").write("\n".join(table) + "\n")
    print("worst cosine vs bf16 oracle", worst)
    x = (img.float() / 255).permute(0, 3, 1, 2)
    logits = ref(x)
    loss = torch.nn.functional.cross_entropy(logits, lab.long())
    loss.backward()
    assert abs(out[0].item() / (N * H * W) - loss.item()) < 3e-2
    for (n, p), (_, q) in zip(own.named_parameters(), ref.named_parameters()):
        if n.endswith((".0.bias", ".3.bias")) and "double_conv" in n:
            continue
        cos = torch.nn.functional.cosine_similarity(p.grad.float().flatten(), q.grad.float().flatten(), dim=0).item()
        assert cos > 0.75, (n, cos)


def test_trainer_cuda_graph_loss_decreases(tmp_path):
    cfg = TrainConfig(base_channels=32, tile=64, train_tiles=16, holdout_tiles=4, batch_size=4, sync_every=1,
                      epochs=3, world_size=1, codec="fp32", device="cuda", out_dir=str(tmp_path),
                      sample_images=2, lr=2e-3)
    tr = Trainer(cfg)
    losses = []
    for _ in range(4):
        losses.append(tr.train_epoch()["mean_loss"])
        tr.epoch += 1
    assert losses[-1] < losses[0]
    ev = tr.evaluate()
    assert ev["eval_loss"] == ev["eval_loss"]
    tr.dump_samples()
    p = tr.save(str(tmp_path / "c.pt"))
    tr2 = Trainer(TrainConfig(**{**cfg.to_dict(), "resume": p, "out_dir": str(tmp_path / "b")}))
    torch.testing.assert_close(tr2.model.flat["param"], tr.model.flat["param"])


def test_engine_bilinear_mode_matches_oracle():
    """The bilinear up-sampling variant (R:609) through the GPU engine against the CPU oracles."""
    torch.manual_seed(1)
    own = UNet(out_classes=6, up_sample_mode="bilinear", base_channels=32).cuda()
    cpu = UNet(out_classes=6, up_sample_mode="bilinear", base_channels=32)
    cpu.load_state_dict(own.state_dict())
    own.flatten_parameters(torch.device("cuda"))
    cpu.flatten_parameters(torch.device("cpu"))
    N, H, W = 4, 128, 128                                  # as above: fewer images make the 4x4-level BNs too noisy
    img = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
    lab = torch.randint(0, 6, (N, H, W), dtype=torch.uint8, device="cuda")
    eng = NativeEngine(own, N, H, W)
    eng.load_batch(img, lab)
    out = eng.train_microbatch().clone()
    eng_cpu = NativeEngine(cpu, N, H, W, act_dtype=torch.bfloat16)
    eng_cpu.load_batch(img.cpu(), lab.cpu())
    out_cpu = eng_cpu.train_microbatch().clone()
    assert abs(out[0].item() - out_cpu[0].item()) / out_cpu[0].item() < 2e-2
    table = []
    for (n, p), (_, q) in zip(own.named_parameters(), cpu.named_parameters()):
        if n.endswith((".0.bias", ".3.bias")) and "double_conv" in n:
            continue
        cos = torch.nn.functional.cosine_similarity(p.grad.float().flatten().cpu(), q.grad.float().flatten(), dim=0).item()
        table.append((cos, n))
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/grad_cosines_bilinear.txt", "w").write("\n".join(f"{c:.4f} {n}" for c, n in table) + "\n")
    # bf16 rounding noise grows faster with depth here than with the transposed conv (smooth up-sampled
    # features, small gradient signal at random init); the CPU oracle with the same rounding points shows
    # the same distance to fp32 autograd layer by layer (0.71 at the stem for both), so only the direction
    # is checked away from the loss
    p_numel = {n: p.numel() for n, p in own.named_parameters()}
    for cos, n in table:
        floor = 0.99 if n.startswith("conv_last") else (0.95 if n.startswith("up_conv1.double_conv") else 0.75)
        if p_numel[n] <= 64 and floor == 0.75:
            floor = 0.5                                  # 32-element BN vectors far from the loss: cosine itself is noisy
        assert cos > floor, (n, cos)


@pytest.mark.parametrize("optimizer", ["adam", "sgd"])
def test_trainer_accumulates_k_microbatches_under_graph(tmp_path, optimizer):
    """The reference's loop shape: K micro-batches accumulate, then one exchange + step (R:756-759)."""
    cfg = TrainConfig(base_channels=32, tile=64, train_tiles=24, holdout_tiles=0, batch_size=2, sync_every=3,
                      epochs=3, world_size=1, codec="fp32", device="cuda", out_dir=str(tmp_path), sample_images=0,
                      optimizer=optimizer, lr=2e-3 if optimizer == "adam" else 5e-2, momentum=0.9)
    tr = Trainer(cfg)
    losses = []
    for _ in range(4):
        st = tr.train_epoch()
        losses.append(st["mean_loss"])
        tr.epoch += 1
    assert st["micro_batches"] == 12
    assert tr.optimizer.step_count == 16                       # 4 epochs x 12 micro-batches / 3
    assert losses[-1] < losses[0]
    assert float(tr.model.flat["grad"].abs().max()) == 0.0     # zeroed by the fused step


def test_engine_headline_shape_tcgen05_vs_cudnn_gradients():
    """The configuration bench.py times (b = 64, 512 x 512; N = 2 here): every gradient tensor of the tcgen05
    path against the SAME engine with the library convolutions (cuDNN through torch), and both against fp32
    eager autograd of the same model (TF32 off) as ground truth.  The two bf16 arms round at the same points
    but sum in different orders; one-ulp differences of bf16 activations are then amplified layer by layer
    (22 BatchNorms), so the direct comparison is tight only near the loss.  The claim that matters and is
    asserted for EVERY tensor: our kernels are as close to the fp32 truth as the library arm is."""
    torch.manual_seed(0)
    a = UNet(out_classes=6, base_channels=64).cuda()
    b = UNet(out_classes=6, base_channels=64).cuda()
    ref = UNet(out_classes=6, base_channels=64).cuda()
    b.load_state_dict(a.state_dict())
    ref.load_state_dict(a.state_dict())
    a.flatten_parameters(torch.device("cuda"))
    b.flatten_parameters(torch.device("cuda"))
    N, H, W = 2, 512, 512
    img = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
    lab = torch.randint(0, 6, (N, H, W), dtype=torch.uint8, device="cuda")
    outs = []
    for model, impl in ((a, "tcgen05"), (b, "cudnn")):
        eng = NativeEngine(model, N, H, W, conv_impl=impl)
        eng.load_batch(img, lab)
        outs.append(eng.train_microbatch().clone())
        del eng
    assert abs(outs[0][0].item() - outs[1][0].item()) <= 2e-3 * abs(outs[1][0].item())
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        loss = torch.nn.functional.cross_entropy(ref((img.float() / 255).permute(0, 3, 1, 2)), lab.long())
        loss.backward()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
    assert abs(outs[0][0].item() / (N * H * W) - loss.item()) < 2e-2
    rows = []
    for (n, p), (_, q), (_, r) in zip(a.named_parameters(), b.named_parameters(), ref.named_parameters()):
        if n.endswith((".0.bias", ".3.bias")) and "double_conv" in n:
            continue                                   # conv biases under BatchNorm: gradient is rounding noise
        ga, gb, gr = p.grad.float(), q.grad.float(), r.grad.float()
        scale = gr.abs().max().item()
        e_tc = (ga - gr).abs().max().item() / max(scale, 1e-12)
        e_lib = (gb - gr).abs().max().item() / max(scale, 1e-12)
        e_ab = (ga - gb).abs().max().item() / max(scale, 1e-12)
        rows.append((n, e_tc, e_lib, e_ab))
        # no further from the truth than the library arm.  Where bf16 rounding noise dominates BOTH arms (far
        # from the loss the library arm itself is 0.25 - 0.8 away from fp32 on the small BN vectors) the two
        # errors are two draws of the same noise and their ratio fluctuates from run to run: a wider band there
        if e_lib < 0.1:
            assert e_tc <= 1.5 * e_lib + 5e-3, (n, e_tc, e_lib)
        else:
            assert e_tc <= 2.5 * e_lib and e_tc < 1.0, (n, e_tc, e_lib)
        if n.startswith("conv_last"):
            assert e_ab <= 1e-2, (n, e_ab)                           # at the loss the two arms agree tightly
    ratios = sorted(e1 / max(e2, 1e-6) for _, e1, e2, _ in rows)
    assert ratios[len(ratios) // 2] <= 1.25, ratios[len(ratios) // 2]      # and on the whole the two arms are equally close
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/grad_errors_headline_shape.txt", "w") as fh:
        fh.write("# max-abs gradient error / max|fp32 gradient|, U-Net b=64, 2 x 512 x 512: tcgen05 arm vs fp32 autograd, "
                 "cuDNN arm vs fp32 autograd, tcgen05 vs cuDNN arm\n")
        for n, e1, e2, e3 in rows:
            fh.write(f"{e1:.4f} {e2:.4f} {e3:.4f} {n}\n")
    print("worst tcgen05-vs-fp32:", max(rows, key=lambda t: t[1]))


def test_convergence_parity_with_reference_arm(reference_script):
    """Same initial state_dict, same synthetic tiles, same order, Adam with the reference's defaults (R:704):
    the reference's own UNet under bf16 autocast (the arm bench.py --impl reference times) against this
    repo's Trainer, 100 optimizer steps each.  "Faster" only counts if it trains to the same place."""
    from baseline.ref_loader import reference_unet
    from distributed_deep_learning_on_personal_computers_b200.utils.data import synthetic_tiles
    torch.manual_seed(0)
    steps, B, T = 100, 4, 128
    tiles = synthetic_tiles(32, T, 6, seed=3)
    cfg = TrainConfig(base_channels=32, tile=T, train_tiles=32, holdout_tiles=0, batch_size=B, sync_every=1,
                      epochs=1, world_size=1, codec="fp32", device="cuda", out_dir="gpurun_out/parity_run",
                      sample_images=0, shuffle=False, cuda_graph=True)
    tr = Trainer(cfg, datasets=(tiles, None))
    ref = reference_unet(32, 6).cuda()
    ref.load_state_dict(tr.model.portable_state_dict())
    ref = ref.to(memory_format=torch.channels_last)
    opt = torch.optim.Adam(ref.parameters())
    crit = torch.nn.CrossEntropyLoss()
    own_loss, ref_loss, own_acc, ref_acc = [], [], [], []
    for s in range(steps):
        i0 = (s * B) % 32
        img, lab = tiles.images[i0:i0 + B].cuda(), tiles.labels[i0:i0 + B].cuda()
        out = tr.train_step(img, lab)
        px = lab.numel()
        own_loss.append(out[0].item() / px)
        own_acc.append(out[1].item() / px)
        x = (img.float() / 255).permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = ref(x)
            loss = crit(logits.float(), lab.long())
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        ref_loss.append(loss.item())
        ref_acc.append((logits.argmax(1) == lab.long()).float().mean().item())
    tail = slice(steps - 20, steps)
    ol, rl = sum(own_loss[tail]) / 20, sum(ref_loss[tail]) / 20
    oa, ra = sum(own_acc[tail]) / 20, sum(ref_acc[tail]) / 20
    import os
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/convergence_parity.txt", "w") as fh:
        fh.write("# step own_loss ref_loss own_acc ref_acc (U-Net b=32, 128x128 tiles, batch 4, Adam lr 1e-3)\n")
        for s in range(0, steps, 5):
            fh.write(f"{s:3d} {own_loss[s]:.4f} {ref_loss[s]:.4f} {own_acc[s]:.4f} {ref_acc[s]:.4f}\n")
        fh.write(f"# mean of last 20 steps: loss own {ol:.4f} ref {rl:.4f}; pixel accuracy own {oa:.4f} ref {ra:.4f}\n")
    print(f"last-20 mean loss own {ol:.4f} ref {rl:.4f}; accuracy own {oa:.4f} ref {ra:.4f}")
    assert own_loss[0] == pytest.approx(ref_loss[0], rel=2e-2)          # same starting point
    assert ol < 0.6 * own_loss[0] and rl < 0.6 * ref_loss[0]            # both actually train
    assert abs(ol - rl) <= 0.03 * rl + 0.01, (ol, rl)                   # same place: within 3 %
    assert abs(oa - ra) <= 0.01, (oa, ra)                               # pixel accuracy within 1 point
