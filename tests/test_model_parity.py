"""Model parity with the reference's own UNet class (Appendix A)."""
import os
import sys

import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200 import UNet
from distributed_deep_learning_on_personal_computers_b200.models.unet import (
    count_parameters, conv_flops_per_tile)

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baseline.ref_loader import load_reference_models, reference_unet  # noqa: E402


@pytest.fixture(scope="module")
def ref_ns(reference_script):
    return load_reference_models(2, reference_script)


def test_state_dict_layout_matches_reference(ref_ns):
    ref = ref_ns.UNet(out_classes=6)
    own = UNet(out_classes=6, base_channels=32)
    rs, os_ = ref.state_dict(), own.state_dict()
    assert list(rs.keys()) == list(os_.keys())
    assert len(rs) == 166
    for k in rs:
        assert rs[k].shape == os_[k].shape and rs[k].dtype == os_[k].dtype, k
    assert [n for n, _ in ref.named_parameters()] == [n for n, _ in own.named_parameters()]
    assert count_parameters(own) == 8_723_558


def test_flops_formula_matches_survey():
    assert abs(conv_flops_per_tile(32, 512) / 1e9 - 113.36) < 0.2


@pytest.mark.parametrize("mode", ["conv_transpose", "bilinear"])
def test_logits_and_grads_match_reference(ref_ns, mode):
    torch.manual_seed(0)
    ns = load_reference_models(8, ref_ns.path)        # width 8: fast on CPU
    ref = ns.UNet(out_classes=6, up_sample_mode=mode)
    own = UNet(out_classes=6, up_sample_mode=mode, base_channels=8)
    own.load_state_dict(ref.state_dict())
    own.flatten_parameters()                          # flat storage must not change the math
    assert own.down_conv1.double_conv.double_conv[0].weight.stride() == (27, 1, 9, 3)
    # compare in float64: at 2x2..4x4 resolution BN turns fp32 rounding into ~1% gradient noise
    ref64 = ns.UNet(out_classes=6, up_sample_mode=mode).double()
    ref64.load_state_dict(ref.state_dict())
    own64 = UNet(out_classes=6, up_sample_mode=mode, base_channels=8).double()
    own64.load_state_dict(own.portable_state_dict())
    x64 = torch.rand(2, 3, 64, 64, dtype=torch.float64)
    y64 = torch.randint(0, 6, (2, 64, 64))
    a, b = ref64(x64), own64(x64)
    torch.testing.assert_close(b, a, rtol=1e-9, atol=1e-10)
    torch.nn.functional.cross_entropy(a, y64).backward()
    torch.nn.functional.cross_entropy(b, y64).backward()
    for (n, p), (_, q) in zip(ref64.named_parameters(), own64.named_parameters()):
        torch.testing.assert_close(q.grad, p.grad, rtol=1e-7, atol=1e-12, msg=n)
    x = torch.rand(2, 3, 64, 64)
    y = torch.randint(0, 6, (2, 64, 64))
    lr_, lo_ = ref(x), own(x)
    torch.testing.assert_close(lo_, lr_, rtol=1e-3, atol=2e-4)
    torch.nn.functional.cross_entropy(lr_, y).backward()
    torch.nn.functional.cross_entropy(lo_, y).backward()
    for (n, p), (_, q) in zip(ref.named_parameters(), own.named_parameters()):
        if not n.startswith(("down_conv1", "up_conv1", "conv_last")):
            continue            # 2x2..8x8 levels: fp32 noise through BN over <100 samples; see f64 above
        # different conv algorithms (flat storage is channels-last) -> fp32 noise that BN amplifies
        scale = p.grad.abs().max().item()
        assert (q.grad - p.grad).abs().max().item() <= 2e-2 * scale + 5e-6, n
    # and the checkpoint goes back into the reference class
    ref2 = ns.UNet(out_classes=6, up_sample_mode=mode)
    ref2.load_state_dict(own.portable_state_dict())


def test_wider_reference_variant(reference_script):
    ref = reference_unet(128)
    own = UNet(out_classes=6, base_channels=128)
    assert count_parameters(ref) == count_parameters(own) == 139_428_230


def test_bad_upsample_mode():
    with pytest.raises(ValueError):
        UNet(up_sample_mode="nearest")


def test_flat_layout_puts_tensor_core_weights_first():
    """The flat buffer starts with the conv / convT weights whose gradients the wgrad kernels write
    (the region the fused exchange reduces in their epilogues); everything small follows."""
    from distributed_deep_learning_on_personal_computers_b200.models.unet import UNet
    torch.manual_seed(0)
    m = UNet(6, "conv_transpose", base_channels=16)
    want = {k: v.detach().clone() for k, v in m.state_dict().items()}
    flat = m.flatten_parameters()
    bulk, offs = flat["bulk_elems"], flat["offsets"]
    assert 0 < bulk < flat["param"].numel() and bulk % 64 == 0
    shapes = dict(m.named_parameters())
    for name, (off, n) in offs.items():
        p = shapes[name]
        is_bulk = off < bulk
        assert (off + n <= bulk) == is_bulk                       # no tensor straddles the boundary
        if is_bulk:
            assert p.dim() == 4 and p.shape[-1] > 1
        else:
            assert p.dim() == 1 or name.startswith("down_conv1.double_conv.double_conv.0") or "conv_last" in name
    small = sum(n for _, (off, n) in offs.items() if off >= bulk)
    assert small < 0.02 * bulk
    for k, v in m.state_dict().items():                           # values and shapes survive the move
        assert torch.equal(v, want[k])
