"""The allocation-free head kernel entry point, called through the C ABI.

Some users of libb200seg.so cannot get device memory when they call into it: a host application may
have given the whole GPU to its own pool and forbid the code it loads from allocating, a
latency-critical loop cannot accept the device-wide stall of cudaMalloc / cudaFree, and a plain C
caller has no framework allocator to borrow a workspace from.  For them ``b200seg_head_ce_static``
keeps its scratch in an array built into the library.  This test drives it exactly as such a caller
would -- ctypes, raw device pointers, a raw stream handle -- and checks it against the ordinary
entry point (``b200seg_head_ce``, caller-provided workspace) on two calls in a row, because the
built-in scratch carries state from one call to the next.

This file sorts last on purpose: if it fails, ``pytest -x`` has already run every other test.
"""
import ctypes

import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200.ops import loader

pytestmark = pytest.mark.gpu

VP, LL, I, F = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_float


def _lib():
    assert loader.library() is not None, loader.load_error()
    lib = ctypes.CDLL(loader.LIB_PATH)
    lib.b200seg_head_ce.restype = I
    lib.b200seg_head_ce.argtypes = [VP, LL, VP, VP, VP, LL, I, I, F, VP, LL, VP, VP, VP, VP, VP, VP]
    lib.b200seg_head_ce_static.restype = I
    lib.b200seg_head_ce_static.argtypes = [VP, LL, VP, VP, VP, LL, I, I, F, VP, LL, VP, VP, VP, VP, VP]
    lib.b200seg_head_ce_workspace_bytes.restype = LL
    lib.b200seg_head_ce_workspace_bytes.argtypes = [I, I]
    lib.b200seg_head_ce_static_scratch_bytes.restype = LL
    return lib


def _outputs(n, h, w, c, k):
    dev = "cuda"
    return dict(da=torch.empty(n, h, w, c, dtype=torch.bfloat16, device=dev),
                dw=torch.zeros(k, c, device=dev), db=torch.zeros(k, device=dev),
                pred=torch.empty(n, h, w, dtype=torch.uint8, device=dev), acc=torch.zeros(2, device=dev))


def test_static_entry_point_matches_workspace_entry_point_twice():
    lib = _lib()
    torch.manual_seed(11)
    n, h, w, c, k = 2, 40, 36, 64, 6
    px = n * h * w
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ws = torch.zeros(lib.b200seg_head_ce_workspace_bytes(c, k), dtype=torch.uint8, device="cuda")
    assert 0 < lib.b200seg_head_ce_static_scratch_bytes() < (1 << 20)
    for call in range(2):                                       # second call sees the first one's leftovers
        a = (torch.randn(n, h, w, c, device="cuda") * (1 + call)).to(torch.bfloat16)
        wt = torch.randn(k, c, device="cuda") * 0.2
        b = torch.randn(k, device="cuda") * 0.1
        lab = torch.randint(0, k, (n, h, w), dtype=torch.uint8, device="cuda")
        ref, got = _outputs(n, h, w, c, k), _outputs(n, h, w, c, k)
        scale = 1.0 / px
        rc = lib.b200seg_head_ce(a.data_ptr(), c, wt.data_ptr(), b.data_ptr(), lab.data_ptr(), px, c, k, scale,
                                 ref["da"].data_ptr(), c, ref["dw"].data_ptr(), ref["db"].data_ptr(),
                                 ref["pred"].data_ptr(), ref["acc"].data_ptr(), ws.data_ptr(), stream)
        assert rc == 0
        rc = lib.b200seg_head_ce_static(a.data_ptr(), c, wt.data_ptr(), b.data_ptr(), lab.data_ptr(), px, c, k,
                                        scale, got["da"].data_ptr(), c, got["dw"].data_ptr(), got["db"].data_ptr(),
                                        got["pred"].data_ptr(), got["acc"].data_ptr(), stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(got["pred"], ref["pred"])
        torch.testing.assert_close(got["da"].float(), ref["da"].float(), rtol=2e-2, atol=1e-6)
        torch.testing.assert_close(got["dw"], ref["dw"], rtol=1e-3, atol=1e-5)
        torch.testing.assert_close(got["db"], ref["db"], rtol=1e-3, atol=1e-6)
        torch.testing.assert_close(got["acc"][0], ref["acc"][0], rtol=1e-4, atol=1e-2)
        assert got["acc"][1].item() == ref["acc"][1].item()


def test_static_entry_point_rejects_sizes_beyond_its_scratch():
    lib = _lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    t = torch.zeros(64, device="cuda")
    for c, k in ((256, 6), (64, 9)):
        rc = lib.b200seg_head_ce_static(t.data_ptr(), c, t.data_ptr(), t.data_ptr(), t.data_ptr(), 32, c, k, 1.0,
                                        None, c, None, None, None, t.data_ptr(), stream)
        assert rc == 1                                          # cudaErrorInvalidValue, nothing launched
    torch.cuda.synchronize()
