"""Gradient codecs: error bounds, the zero guard, bit parity with the reference's arithmetic."""
import pytest
import torch
from hypothesis import example, given, settings, strategies as st

from distributed_deep_learning_on_personal_computers_b200.parallel.codecs import (
    get_codec, reference_restatement)

finite = st.floats(min_value=-1e3, max_value=1e3, allow_nan=False, width=32)


@settings(max_examples=50, deadline=None, derandomize=True)
@given(st.lists(finite, min_size=1, max_size=300), st.sampled_from(["q201", "q21", "int8s", "bf16", "fp16", "fp32"]))
@example([1e-05], "fp16")            # fp16 subnormal: the bound is the subnormal spacing, not relative
def test_roundtrip_error_bound(vals, name):
    g = torch.tensor(vals, dtype=torch.float32)
    c = get_codec(name)
    q, s = c.encode(g)
    back = c.decode(q, s)
    mx = g.abs().max().item()
    bound = c.max_abs_error(mx if c.needs_scale else max(mx, 1e-30))
    assert (back - g).abs().max().item() <= bound * 1.001 + 1e-12


@pytest.mark.parametrize("name,mode", [("q201", "float16"), ("q21", "int8")])
def test_bit_parity_with_reference_arithmetic(name, mode):
    torch.manual_seed(0)
    g = torch.randn(10_000) * 0.01
    c = get_codec(name)
    q, s = c.encode(g)
    assert torch.equal(c.decode(q, s), reference_restatement(g, mode))
    assert q.dtype == (torch.float16 if mode == "float16" else torch.int8)
    assert q.float().abs().max() <= (100 if mode == "float16" else 10)


def test_reference_names_are_accepted():
    assert get_codec("float16").name == "q201"
    assert get_codec("int8").name == "q21"           # the reference's int8 mode (R:474)
    assert get_codec("int8s").levels == 127.0
    assert get_codec("float32").name == "fp32"


@pytest.mark.parametrize("name", ["q201", "q21", "int8s", "fp8", "bf16", "fp32"])
def test_zero_gradient_does_not_crash(name):
    c = get_codec(name)
    g = torch.zeros(64)
    q, s = c.encode(g)
    assert torch.equal(c.decode(q, s), g)


def test_decode_accumulate():
    c = get_codec("q201")
    g = torch.randn(100)
    q, s = c.encode(g)
    out = torch.ones(100)
    c.decode(q, s, out=out, accumulate=True)
    torch.testing.assert_close(out, 1 + c.decode(q, s))


def test_unknown_codec():
    with pytest.raises(ValueError):
        get_codec("zip")
