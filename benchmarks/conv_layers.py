#!/usr/bin/env python
"""Per-layer conv timing: our tcgen05 kernels vs cuDNN (through torch) on the flagship shapes.

    python benchmarks/conv_layers.py [--base 64] [--batch 16] [--tile 512] [--reps 5] [--out FILE]

Device time with CUDA events, 3 warm-up calls, inputs of the big layers are >> L2.
Prints one JSON line per (layer, op): ms for ours / cuDNN, TFLOP/s, fraction of the measured bf16 peak.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn  # noqa: E402


def timeit(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--base", type=int, default=64)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--tile", type=int, default=512)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--ops", default="fprop,dgrad,wgrad")
    ap.add_argument("--only", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    peak = 1405.9
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        pass
    b, N, T = args.base, args.batch, args.tile
    w = [b, 2 * b, 4 * b, 8 * b, 8 * b]
    layers = [("L0.stem", 0, 16, w[0]), ("L0.c1", 0, w[0], w[0]), ("L0.dec0", 0, w[1] + w[0], w[0]),
              ("L1.c0", 1, w[0], w[1]), ("L1.c1", 1, w[1], w[1]), ("L1.dec0", 1, w[2] + w[1], w[1]),
              ("L2.c0", 2, w[1], w[2]), ("L2.c1", 2, w[2], w[2]), ("L2.dec0", 2, w[3] + w[2], w[2]),
              ("L3.c0", 3, w[2], w[3]), ("L3.c1", 3, w[3], w[3]), ("L3.dec0", 3, w[3] + w[3], w[3]),
              ("L4.c0", 4, w[3], w[4]), ("L4.dec0", 4, w[4] + w[4], w[3]), ("L5.c0", 5, w[4], w[4])]
    if args.only:
        layers = [l for l in layers if l[0] in args.only.split(",")]
    rows = []
    for name, lvl, cin, cout in layers:
        H = W = T >> lvl
        x = torch.randn(N, H, W, cin, device="cuda").to(torch.bfloat16)
        dy = torch.randn(N, H, W, cout, device="cuda").to(torch.bfloat16)
        wt = (torch.randn(cout, 3, 3, cin, device="cuda") * 0.05).to(torch.bfloat16)
        bias = torch.zeros(cout, device="cuda")
        y = torch.empty(N, H, W, cout, dtype=torch.bfloat16, device="cuda")
        dx = torch.empty(N, H, W, cin, dtype=torch.bfloat16, device="cuda")
        dw = torch.zeros(cout, 3, 3, cin, device="cuda")
        flops = 2.0 * N * H * W * 9 * cin * cout
        xn, wn, dyn = x.permute(0, 3, 1, 2), wt.permute(0, 3, 1, 2), dy.permute(0, 3, 1, 2)
        # cuDNN backward baselines go through autograd of F.conv2d on channels-last tensors, i.e. the very
        # call the reference arm makes (R:579 under loss.backward()): conv2d_input on permuted views
        # landed on a slow cuDNN plan and made dgrad look 3x worse than it is inside a real step.
        xg = xn.detach().requires_grad_(True)
        wg = wn.detach().contiguous(memory_format=torch.channels_last).requires_grad_(True)
        yg = torch.nn.functional.conv2d(xg, wg, None, padding=1)

        def lib_grad(inp):
            def run():
                torch.autograd.grad(yg, inp, dyn, retain_graph=True)
            return run
        ops = {
            # no conv bias, as in the engine (BatchNorm absorbs it, see bn_finalize)
            "fprop": (lambda: Fn.conv3x3_fwd(x, wt, None, y, True, impl="tcgen05"),
                      lambda: torch.nn.functional.conv2d(xn, wn, None, padding=1)),
            "dgrad": (lambda: Fn.conv3x3_dgrad(dy, wt, dx, impl="tcgen05"),
                      lib_grad(xg)),
            "wgrad": (lambda: Fn.conv3x3_wgrad(x, dy, dw, impl="tcgen05"),
                      lib_grad(wg)),
        }
        torch.backends.cudnn.benchmark = True
        for op in args.ops.split(","):
            if op == "dgrad" and cin == 16:
                continue
            ours, lib = ops[op]
            t_o, t_l = timeit(ours, args.reps), timeit(lib, args.reps)
            row = {"layer": name, "op": op, "HW": H, "cin": cin, "cout": cout, "ours_ms": round(t_o, 4),
                   "cudnn_ms": round(t_l, 4), "ours_tflops": round(flops / t_o / 1e9, 1),
                   "cudnn_tflops": round(flops / t_l / 1e9, 1), "ours_frac_of_measured_peak": round(flops / t_o / 1e9 / peak, 3)}
            rows.append(row)
            print(json.dumps(row), flush=True)
        del x, dy, y, dx, xg, wg, yg
    tot = {}
    for r in rows:
        t = tot.setdefault(r["op"], [0.0, 0.0])
        t[0] += r["ours_ms"]
        t[1] += r["cudnn_ms"]
    print(json.dumps({"totals_ms": {k: {"ours": round(v[0], 3), "cudnn": round(v[1], 3)} for k, v in tot.items()}}))
    if args.out:
        with open(args.out, "w") as fh:
            for r in rows:
                fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
