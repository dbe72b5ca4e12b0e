#!/usr/bin/env python
"""Summarise ``ncu --metrics gpu__time_duration.sum --csv`` of one training step.

    python benchmarks/launches_summarize.py gpurun_out/launches4.csv profiles/launches_one_step.txt

Writes the per-kernel totals and then every launch in order (duration, grid, block, name).
"""
import collections
import csv
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    launches = []
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        launches.append((us, r["Grid Size"], r["Block Size"], r["Kernel Name"]))
    total = sum(l[0] for l in launches)
    agg = collections.OrderedDict()
    for us, _, _, name in launches:
        key = name.split("(")[0][:80]
        a = agg.setdefault(key, [0.0, 0])
        a[0] += us
        a[1] += 1
    out = ["# every kernel launch of ONE training step (fwd + loss + bwd + Adam), U-Net b=64, batch 16, 512x512, 1 GPU",
           "# ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none (serialised, caches flushed",
           "# between launches: compare SHARES, not absolute times)",
           f"# launches {len(launches)}, total {total / 1e3:.3f} ms", "", "## totals per kernel"]
    for key, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        out.append(f"{us / 1e3:9.3f} ms {100 * us / total:5.1f}%  x{n:4d}  {key}")
    out += ["", "## launches in order (us, grid, block, kernel)"]
    for i, (us, grid, block, name) in enumerate(launches):
        out.append(f"{i:4d} {us:9.1f}  {grid:>14s} {block:>12s}  {name.split('(')[0][:90]}")
    open(dst, "w").write("\n".join(out) + "\n")
    print(f"{len(launches)} launches, {total / 1e3:.3f} ms -> {dst}")


if __name__ == "__main__":
    main()
