"""The explicit forward/backward plan against eager autograd of the same model (CPU, fp32)."""
import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200 import UNet
from distributed_deep_learning_on_personal_computers_b200.models.engine import NativeEngine


def _run(mode, n=2, h=64, w=64, b=8, fused_tail=False):
    torch.manual_seed(1)
    own = UNet(out_classes=6, base_channels=b, up_sample_mode=mode)
    ref = UNet(out_classes=6, base_channels=b, up_sample_mode=mode)
    ref.load_state_dict(own.state_dict())
    own.flatten_parameters()
    img = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8)
    lab = torch.randint(0, 6, (n, h, w), dtype=torch.uint8)
    eng = NativeEngine(own, n, h, w, fused_tail=fused_tail)
    assert eng.fused_tail == fused_tail
    eng.load_batch(img, lab)
    out = eng.train_microbatch().clone()
    x = (img.float() / 255).permute(0, 3, 1, 2).contiguous()
    logits = ref(x)
    loss = torch.nn.functional.cross_entropy(logits, lab.long())
    loss.backward()
    return own, ref, eng, out, logits, loss, lab


@pytest.mark.parametrize("fused_tail", [False, True])     # True: last BN+ReLU, head and BN-backward sums as one op
@pytest.mark.parametrize("mode", ["conv_transpose", "bilinear"])
def test_engine_matches_autograd(mode, fused_tail):
    own, ref, eng, out, logits, loss, lab = _run(mode, fused_tail=fused_tail)
    px = lab.numel()
    assert abs(out[0].item() / px - loss.item()) < 1e-5
    assert out[1].item() == (logits.argmax(1) == lab.long()).sum().item()
    for (n, p), (_, q) in zip(own.named_parameters(), ref.named_parameters()):
        scale = q.grad.abs().max().item()
        # conv biases that feed a BatchNorm have a mathematically zero gradient: noise only
        atol = 1e-6 if (".double_conv." in n and n.endswith((".0.bias", ".3.bias"))) else 2e-2 * scale + 1e-9
        assert (p.grad - q.grad).abs().max().item() <= atol + 2e-4 * scale, n
    for (n, a), (_, b) in zip(own.named_buffers(), ref.named_buffers()):
        torch.testing.assert_close(a.float(), b.float(), rtol=1e-4, atol=1e-5, msg=n)


def test_gradients_accumulate_across_microbatches():
    own, ref, eng, *_ = _run("conv_transpose")
    g1 = own.flat["grad"].clone()
    eng.train_microbatch()
    # second pass over the same batch: BN batch statistics are identical, so grads double
    torch.testing.assert_close(own.flat["grad"], 2 * g1, rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("fused_tail", [False, True])
def test_eval_uses_running_statistics(fused_tail):
    own, ref, eng, *_ = _run("conv_transpose", fused_tail=fused_tail)
    ref.eval()
    img = torch.randint(0, 256, (2, 64, 64, 3), dtype=torch.uint8)
    lab = torch.randint(0, 6, (2, 64, 64), dtype=torch.uint8)
    eng.load_batch(img, lab)
    before = [b.clone() for b in own.buffers()]
    out = eng.evaluate().clone()
    for a, b in zip(own.buffers(), before):
        assert torch.equal(a, b)                         # eval must not touch the buffers
    with torch.no_grad():
        logits = ref((img.float() / 255).permute(0, 3, 1, 2))
    want = torch.nn.functional.cross_entropy(logits, lab.long(), reduction="sum")
    assert abs(out[0].item() - want.item()) / want.item() < 1e-4
    assert torch.equal(eng.pred.long(), logits.argmax(1))


def test_rejects_bad_tile_size():
    m = UNet(base_channels=8)
    m.flatten_parameters()
    with pytest.raises(ValueError):
        NativeEngine(m, 1, 48, 64)
