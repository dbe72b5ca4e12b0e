"""End-to-end on the GPU: native engine vs eager autograd, CUDA-graph step, loss decreases."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig, UNet
from distributed_deep_learning_on_personal_computers_b200.models.engine import NativeEngine
from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("conv", ["cudnn", "tcgen05"])
def test_engine_gradients_match_oracle(conv):
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
    eng = NativeEngine(own, N, H, W, conv_impl=conv)
    eng.load_batch(img, lab)
    out = eng.train_microbatch().clone()
    eng_cpu = NativeEngine(cpu, N, H, W, act_dtype=torch.bfloat16)
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
    open(f"gpurun_out/grad_cosines_{conv}.txt", "w").write("\n".join(table) + "\n")
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
