#!/usr/bin/env python
"""Turn an ``.ncu-rep`` into the tab-separated summary that is committed under ``profiles/``.

    python benchmarks/ncu_summarize.py gpurun_out/prof3_conv.ncu-rep profiles/ncu_conv_kernels.tsv "title line"

One row per captured launch; the metrics are the ones the roofline discussion in DESIGN.md uses.
"""
import csv
import io
import subprocess
import sys

METRICS = [
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else rep
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    cols = [m for m in METRICS if m in idx]
    with open(out, "w") as f:
        f.write(f"# {title}\n")
        f.write("# ncu --set full --clock-control none (B200, sm_100a); one row per captured launch; units: " +
                ", ".join(f"{m}[{units[idx[m]]}]" for m in cols[3:]) + "\n")
        f.write("\t".join(["Kernel Name"] + cols) + "\n")
        for r in body:
            name = r[idx["Kernel Name"]][:70]
            f.write("\t".join([name] + [r[idx[m]] for m in cols]) + "\n")
    print(f"{len(body)} launches -> {out}")


if __name__ == "__main__":
    main()
