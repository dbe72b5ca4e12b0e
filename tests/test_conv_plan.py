"""Which tcgen05 kernel variant each benchmarked layer shape takes (host-side planner, no GPU needed).

The fast paths are selected by shape rules in ``csrc/conv/conv_tc.cu``; a rule that silently stops matching would
leave a layer on a slower variant with every numerics test still green.  ``b200conv_plan`` runs the same host code
as a launch (dry) and reports the decision, so the table the per-layer timings were measured with
(``profiles/conv_plan_b64.txt``) is pinned here."""
import os

import pytest

from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn, loader

pytestmark = pytest.mark.skipif(loader.library() is None, reason=f"native library not built: {loader.load_error()}")
SMEM_LIMIT = 227 * 1024
N = 16


def _no_switches():
    return not any(k.startswith("B200SEG_") and k not in ("B200SEG_LIB",) for k in os.environ)


def _plan(*a):
    return Fn.conv_plan(*a)


@pytest.mark.skipif(not _no_switches(), reason="B200SEG_* switches change the plan on purpose")
def test_headline_layers_take_the_measured_variants():
    # 64-channel 512 x 512 layers: halo strips, resident filter bank, four-row N-merged units, warp-private epilogue
    for op in ("fprop", "dgrad"):
        p = _plan(op, N, 512, 512, 64, 64)
        assert (p["kernel"], p["block_n"], p["rows_per_unit"], p["weights_resident"], p["private_epilogue"]) == \
            ("tap_gemm/strip", 64, 4, 1, 1), (op, p)
    p = _plan("fprop", N, 512, 512, 16, 64)                      # stem: 3 (padded to 16) -> 64
    assert (p["kernel"], p["block_k"], p["rows_per_unit"], p["weights_resident"]) == ("tap_gemm/strip", 16, 4, 1)
    p = _plan("fprop", N, 512, 512, 192, 64)                     # decoder's first conv: streamed weights, two-row units
    assert (p["kernel"], p["rows_per_unit"], p["weights_resident"], p["private_epilogue"]) == ("tap_gemm/strip", 2, 0, 1)
    p = _plan("dgrad", N, 512, 512, 192, 64)                     # its dgrad: three resident 64-column N tiles
    assert (p["block_n"], p["n_tiles"], p["rows_per_unit"], p["weights_resident"]) == (64, 3, 4, 1)
    p = _plan("fprop", N, 256, 256, 64, 128)                     # 64 -> 128: two resident 64-column N tiles
    assert (p["block_n"], p["n_tiles"], p["rows_per_unit"], p["weights_resident"]) == (64, 2, 4, 1)
    p = _plan("fprop", N, 256, 256, 128, 128)                    # 128-column tiles, two-row units
    assert (p["kernel"], p["block_n"], p["rows_per_unit"]) == ("tap_gemm/strip", 128, 2)
    p = _plan("fprop", N, 128, 128, 256, 256)                    # 256-channel strip layers: two 128-column N tiles
    assert (p["kernel"], p["block_n"], p["n_tiles"], p["rows_per_unit"]) == ("tap_gemm/strip", 128, 2, 2)
    for hw, c in ((64, 512), (32, 512), (16, 512)):              # rows shorter than 128 pixels: square pixel boxes
        assert _plan("fprop", N, hw, hw, c, c)["kernel"] == "tap_gemm/box"
    p = _plan("fprop", N, 16, 16, 512, 512)                      # bottleneck: 32 pixel tiles -> 128-column N tiles (128 CTAs, not 64)
    assert (p["block_n"], p["n_tiles"], p["grid"]) == (128, 4, 128)
    assert _plan("fprop", N, 32, 32, 512, 512)["block_n"] == 256  # enough tiles: 256-column N tiles stay
    # weight gradient: row pairs where rows are >= 64 px and Cin <= 512, split-K below / beyond
    assert _plan("wgrad", N, 512, 512, 64, 64)["kernel"] == "wgrad/row-pair"
    assert _plan("wgrad", N, 512, 512, 192, 64)["kernel"] == "wgrad/row-pair"
    assert _plan("wgrad", N, 64, 64, 512, 512)["kernel"] == "wgrad/row-pair"
    assert _plan("wgrad", N, 64, 64, 1024, 512)["kernel"] == "wgrad/split-k"
    assert _plan("wgrad", N, 32, 32, 512, 512)["kernel"] == "wgrad/split-k"


def test_every_unet_layer_fits_the_machine():
    from distributed_deep_learning_on_personal_computers_b200 import UNet
    for base in (32, 64, 128):
        m = UNet(out_classes=6, base_channels=base)
        res = {f"down_conv{i}": 512 >> (i - 1) for i in range(1, 6)}
        res.update({"double_conv": 16}, **{f"up_conv{i}": 512 >> (i - 1) for i in range(1, 6)})
        for block, pair in m.conv_specs().items():
            for sp in pair:
                hw, cin = res[block], max(sp.cin, 16)
                for op in ("fprop", "dgrad"):
                    if op == "dgrad" and sp.cin < 16:
                        continue                                  # the stem needs no data gradient
                    p = _plan(op, N, hw, hw, cin, sp.cout)
                    assert p["smem_bytes"] <= SMEM_LIMIT and 1 <= p["grid"] <= 148 and p["stages"] >= 2, (sp.name, op, p)
                    assert p["block_n"] * p["n_tiles"] >= (sp.cout if op == "fprop" else cin)
                    if p["kernel"] == "tap_gemm/strip":
                        assert hw % 128 == 0 and p["strip_slots"] >= 4
                assert _plan("wgrad", N, hw, hw, cin, sp.cout)["kernel"].startswith("wgrad/")


def test_plan_rejects_unsupported_channel_counts():
    with pytest.raises(RuntimeError):
        _plan("fprop", 1, 64, 64, 24, 20)                        # Cout not a multiple of 16
