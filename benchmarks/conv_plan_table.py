#!/usr/bin/env python
"""Print which tcgen05 kernel variant every distinct conv layer shape of a U-Net config takes (host-only, no GPU):

    python benchmarks/conv_plan_table.py --base 64 --batch 16 --tile 512 > profiles/conv_plan_b64.txt
"""
import argparse
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_deep_learning_on_personal_computers_b200.ops import functional as Fn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--base", type=int, default=64)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--tile", type=int, default=512)
    a = ap.parse_args()
    b, n, t = a.base, a.batch, a.tile
    w = [b, 2 * b, 4 * b, 8 * b, 8 * b]
    layers = [("L0.stem", 0, 16, w[0]), ("L0.c1", 0, w[0], w[0]), ("L0.dec0", 0, w[1] + w[0], w[0]),
              ("L1.c0", 1, w[0], w[1]), ("L1.c1", 1, w[1], w[1]), ("L1.dec0", 1, w[2] + w[1], w[1]),
              ("L2.c0", 2, w[1], w[2]), ("L2.c1", 2, w[2], w[2]), ("L2.dec0", 2, w[3] + w[2], w[2]),
              ("L3.c0", 3, w[2], w[3]), ("L3.c1", 3, w[3], w[3]), ("L3.dec0", 3, w[3] + w[3], w[3]),
              ("L4.c0", 4, w[3], w[4]), ("L4.dec0", 4, w[4] + w[4], w[3]), ("L5.c0", 5, w[4], w[4])]
    head = subprocess.run(["git", "-C", os.path.dirname(os.path.abspath(__file__)), "rev-parse", "--short", "HEAD"],
                          capture_output=True, text=True).stdout.strip()
    sw = {k: v for k, v in os.environ.items() if k.startswith("B200SEG_")}
    print(f"# kernel variant per conv layer, U-Net b={b}, batch {n}, {t}x{t}; from b200conv_plan (the launchers' own host code, dry);"
          f" commit {head}; switches in effect: {sw or 'none'}")
    print("# rows = output rows per work unit (4: vertical taps merged along N over a resident filter bank; 2: streamed weights);"
          " res = filter bank resident in shared memory; priv = warp-private epilogue")
    print(f"{'layer':9s} {'op':6s} {'HW':>4s} {'cin':>5s} {'cout':>5s}  {'kernel':16s} {'BK':>3s} {'BN':>4s} {'Ntiles':>6s} {'rows':>4s} {'res':>3s} "
          f"{'priv':>4s} {'strips':>6s} {'stages':>6s} {'grid':>4s} {'smem KB':>8s}")
    for name, lvl, cin, cout in layers:
        hw = t >> lvl
        for op in ("fprop", "dgrad", "wgrad"):
            if op == "dgrad" and name == "L0.stem":
                continue
            p = Fn.conv_plan(op, n, hw, hw, cin, cout)
            if op == "wgrad":
                print(f"{name:9s} {op:6s} {hw:4d} {cin:5d} {cout:5d}  {p['kernel']:16s}")
                continue
            print(f"{name:9s} {op:6s} {hw:4d} {cin:5d} {cout:5d}  {p['kernel']:16s} {p['block_k']:3d} {p['block_n']:4d} {p['n_tiles']:6d} "
                  f"{p['rows_per_unit']:4d} {p['weights_resident']:3d} {p['private_epilogue']:4d} {p['strip_slots']:6d} {p['stages']:6d} "
                  f"{p['grid']:4d} {p['smem_bytes'] / 1024:8.1f}")
    for lvl, c in ((4, w[4]), (3, w[3]), (2, w[3]), (1, w[2]), (0, w[1])):      # transposed convs (input level, channels)
        hw = t >> (lvl + 1)
        for op in ("convT_fwd", "convT_dgrad"):
            p = Fn.conv_plan(op, n, hw, hw, c, c)
            print(f"{'up' + str(lvl + 1):9s} {op[6:]:6s} {hw:4d} {c:5d} {c:5d}  {p['kernel']:16s} {p['block_k']:3d} {p['block_n']:4d} {p['n_tiles']:6d} "
                  f"{p['rows_per_unit']:4d} {p['weights_resident']:3d} {p['private_epilogue']:4d} {p['strip_slots']:6d} {p['stages']:6d} "
                  f"{p['grid']:4d} {p['smem_bytes'] / 1024:8.1f}")


if __name__ == "__main__":
    main()
