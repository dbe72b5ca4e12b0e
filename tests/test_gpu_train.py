"""End-to-end on the GPU: native engine vs eager autograd, CUDA-graph step, loss decreases."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig, UNet
from distributed_deep_learning_on_personal_computers_b200.models.engine import NativeEngine
from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("conv", ["cudnn", "tcgen05"])
def test_engine_gradients_match_autograd(conv):
    torch.manual_seed(0)
    own = UNet(out_classes=6, base_channels=32).cuda()
    ref = UNet(out_classes=6, base_channels=32).cuda()
    ref.load_state_dict(own.state_dict())
    own.flatten_parameters(torch.device("cuda"))
    N, H, W = 2, 64, 64
    img = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda")
    lab = torch.randint(0, 6, (N, H, W), dtype=torch.uint8, device="cuda")
    eng = NativeEngine(own, N, H, W, conv_impl=conv)
    eng.load_batch(img, lab)
    out = eng.train_microbatch().clone()
    x = (img.float() / 255).permute(0, 3, 1, 2)
    logits = ref(x)
    loss = torch.nn.functional.cross_entropy(logits, lab.long())
    loss.backward()
    assert abs(out[0].item() / (N * H * W) - loss.item()) < 3e-2
    # bf16 activations vs fp32 autograd: compare directions, layer by layer
    worst = 1.0
    for (n, p), (_, q) in zip(own.named_parameters(), ref.named_parameters()):
        if n.endswith((".0.bias", ".3.bias")) and "double_conv" in n:
            continue
        a, b = p.grad.float().flatten(), q.grad.float().flatten()
        cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
        worst = min(worst, cos)
        assert cos > 0.97, (n, cos)
    print("worst cosine", worst)


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
