#!/bin/bash
# ncu --set full with source for the one-pass tail kernel (one launch), after a plain run of the same program
mkdir -p gpurun_out
python benchmarks/tail_bench.py > gpurun_out/t7_plain.json 2>&1 || { echo plain failed; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"head_tail_kernel" -s 3 -c 1 -o gpurun_out/prof_tail python benchmarks/tail_bench.py > gpurun_out/t7_ncu.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/t7_ncu.log
ls -la gpurun_out/prof_tail.ncu-rep
