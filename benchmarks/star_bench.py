#!/usr/bin/env python
"""The TCP-star transport's byte movers on a gradient-sized message (host only, no GPU): native library vs the Python
implementation of the same protocol -- deflate / inflate of a q201 (fp16 levels) gradient of the b = 64 model, and a
framed send / receive over a local socket.  ``python benchmarks/star_bench.py > profiles/star_transport.txt``"""
import os
import socket
import statistics
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_deep_learning_on_personal_computers_b200.parallel import star  # noqa: E402


def med(fn, reps=7):
    ts = []
    out = None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append(1e3 * (time.perf_counter() - t0))
    return statistics.median(ts), min(ts), out


def main():
    n = 34_869_446                                           # parameters of the b = 64 U-Net
    rng = np.random.default_rng(0)
    g = np.zeros(n, dtype=np.float16)                        # float16 mode: 201 levels, ~94 % zeros on a real gradient
    idx = rng.choice(n, n // 16, replace=False)
    g[idx] = np.round(rng.standard_normal(idx.size) * 10).astype(np.float16)
    data = g.tobytes()
    native = star._native
    threads = min(12, os.cpu_count() or 1)
    print(f"# TCP-star byte movers, {len(data) / 1e6:.1f} MB message (q201 gradient of the b=64 model, 1/16 non-zero), "
          f"{os.cpu_count()} host cores, {threads} threads, 1 MB blocks, level 1; median (min) of 7, ms")
    print(f"# native = csrc/host/star_native.cpp (lib/libb200star.so), python = parallel/star.py fallback; native available: {native is not None}")
    print(f"{'implementation':15s} {'strategy':9s} {'deflate ms':>16s} {'inflate ms':>16s} {'wire MB':>8s} {'ratio':>6s}")
    for impl in ("native", "python"):
        if impl == "native" and native is None:
            continue
        star._native = native if impl == "native" else None
        for strat in ("default", "rle"):
            dm, dmin, z = med(lambda: star.deflate(data, 1, threads, 1_000_000, strat))
            im, imin, o = med(lambda: star.inflate(z, threads))
            assert o == data
            print(f"{impl:15s} {strat:9s} {dm:9.0f} ({dmin:4.0f}) {im:9.0f} ({imin:4.0f}) {len(z) / 1e6:8.2f} {len(data) / len(z):6.1f}")
    payload = star.deflate(data, 1, threads, 1_000_000, "rle") * 8        # ~37 MB on the wire
    for impl in ("native", "python"):
        if impl == "native" and native is None:
            continue
        star._native = native if impl == "native" else None

        def once():
            a, b = socket.socketpair()
            a.settimeout(30)
            b.settimeout(30)
            th = threading.Thread(target=star.send_frame, args=(a, payload))
            th.start()
            got = star.recv_frame(b)
            th.join()
            a.close()
            b.close()
            return len(got)
        m, mn, k = med(once)
        assert k == len(payload)
        print(f"{impl:15s} frame of {len(payload) / 1e6:.1f} MB over a local socket pair: {m:.0f} ({mn:.0f}) ms = {len(payload) / 1e6 / m:.1f} GB/s")
    star._native = native


if __name__ == "__main__":
    main()
