"""Finds, builds and loads the in-tree native library ``lib/libb200seg.so``.

The library holds every sm_100a kernel plus their ``extern "C"`` launchers (``include/b200seg.h``)
and a ``TORCH_LIBRARY`` binding (``csrc/bindings.cpp``), so one file serves ``torch.ops.b200seg.*``
and plain C callers.  It is built in-tree (``build()``; ninja through ``torch.utils.cpp_extension``)
with ``-gencode arch=compute_100a,code=sm_100a`` and shipped to GPU boxes as a file; importing the
package never compiles anything.
"""
from __future__ import annotations

import glob
import os
import shutil
from typing import Optional

import torch

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPO_DIR = os.path.dirname(PKG_DIR)
CSRC_DIR = os.path.join(REPO_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.environ.get("B200SEG_LIB") or os.path.join(LIB_DIR, "libb200seg.so")   # B200SEG_LIB: A/B another build

_lib_loaded: Optional[str] = None
_load_error: Optional[str] = None
_tried = False

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17", "--use_fast_math",
    "--expt-relaxed-constexpr", "-Xptxas", "-v",
    "-U__CUDA_NO_HALF_OPERATORS__", "-U__CUDA_NO_HALF_CONVERSIONS__",
    "-U__CUDA_NO_BFLOAT16_CONVERSIONS__", "-U__CUDA_NO_HALF2_OPERATORS__",
]


def sources() -> list:
    out = []
    for pat in ("*.cpp", "*.cu", "*/*.cpp", "*/*.cu"):
        out += glob.glob(os.path.join(CSRC_DIR, pat))
    # csrc/host/ is plain C++ for the host-side library (build_host), not part of the CUDA extension
    return sorted(s for s in out if not os.path.basename(s).startswith("standalone_")
                  and os.path.basename(os.path.dirname(s)) != "host")


HOST_LIB_PATH = os.path.join(LIB_DIR, "libb200star.so")


def build_host(verbose: bool = False) -> Optional[str]:
    """Compile the host-side transport (``csrc/host/star_native.cpp``: parallel gzip, framed socket I/O, concurrent
    gather) into ``lib/libb200star.so`` with the system compiler.  Returns the path, or None when it cannot be built
    (no g++ / zlib headers): ``parallel.star`` then runs its Python implementation of the same protocol."""
    import subprocess
    src = os.path.join(CSRC_DIR, "host", "star_native.cpp")
    os.makedirs(LIB_DIR, exist_ok=True)
    tmp = HOST_LIB_PATH + ".tmp"
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-shared", "-fPIC", src, "-o", tmp, "-lz", "-lpthread"]
    try:
        res = subprocess.run(cmd, capture_output=True, text=True)
    except OSError as exc:
        print(f"[b200seg] host library not built: {exc}")
        return None
    if res.returncode != 0:
        print(f"[b200seg] host library not built:\n{res.stderr[-2000:]}")
        return None
    if verbose and res.stderr:
        print(res.stderr)
    os.replace(tmp, HOST_LIB_PATH)
    return HOST_LIB_PATH


def build(verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a and link ``lib/libb200seg.so`` (no GPU needed)."""
    global _tried, _lib_loaded
    build_host(verbose)
    if _lib_loaded is not None:
        # cpp_extension.load() also LOADS what it built, and a second copy of the same TORCH_LIBRARY in one process
        # aborts: with the library already in, compile in a child process and keep running the loaded one.
        import subprocess
        import sys
        code = ("import sys; sys.path.insert(0, %r); "
                "from distributed_deep_learning_on_personal_computers_b200.ops import loader; "
                "loader.build(verbose=%r)" % (REPO_DIR, bool(verbose)))
        subprocess.run([sys.executable, "-c", code], check=True)
        return LIB_PATH
    from torch.utils import cpp_extension
    os.makedirs(LIB_DIR, exist_ok=True)
    build_dir = os.path.join(REPO_DIR, "build", "b200seg")
    os.makedirs(build_dir, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    cpp_extension.load(
        name="b200seg_C",
        sources=sources(),
        extra_include_paths=[os.path.join(REPO_DIR, "include"), CSRC_DIR],
        extra_cflags=["-O3", "-std=c++17"],
        extra_cuda_cflags=NVCC_FLAGS,
        extra_ldflags=[],
        build_directory=build_dir,
        is_python_module=False,
        verbose=verbose,
    )
    built = os.path.join(build_dir, "b200seg_C.so")
    shutil.copy2(built, LIB_PATH)
    # load() registered torch.ops.b200seg from the build directory's file in THIS process; lib/libb200seg.so is a
    # copy of it and must not be loaded on top (double registration of a TORCH_LIBRARY aborts the interpreter)
    _lib_loaded, _tried = built, True
    return LIB_PATH


def library() -> Optional[str]:
    """Path of the loaded library, loading it on first use; ``None`` if it is not built."""
    global _lib_loaded, _load_error, _tried
    if _lib_loaded or _tried:
        return _lib_loaded
    _tried = True
    if not os.path.exists(LIB_PATH):
        _load_error = f"{LIB_PATH} does not exist"
        return None
    try:
        torch.ops.load_library(LIB_PATH)
        _lib_loaded = LIB_PATH
    except Exception as exc:  # pragma: no cover - depends on the box
        _load_error = f"{type(exc).__name__}: {exc}"
    return _lib_loaded


def load_error() -> Optional[str]:
    return _load_error


_error_record = None


def arm_error_record() -> None:
    """Register a pinned host buffer as the tcgen05 kernels' error record on the current device (idempotent)."""
    global _error_record
    if _error_record is None and library() is not None and torch.cuda.is_available():
        _error_record = torch.zeros(8, dtype=torch.int32).pin_memory()
        torch.ops.b200seg.set_error_record(_error_record)


def last_device_error() -> Optional[str]:
    """What a bounded mbarrier wait left behind before it trapped (readable without any CUDA call), or None."""
    r = _error_record
    if r is None or (int(r[0]) & 0xFFFFFFFF) != 0xB200DEAD:
        return None
    return (f"tcgen05 kernel: mbarrier at shared address 0x{int(r[1]) & 0xFFFFFFFF:x} (parity {int(r[2])}) did not flip; "
            f"CTA {int(r[3])}, thread {int(r[4])}")
