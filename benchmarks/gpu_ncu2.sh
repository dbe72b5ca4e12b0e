#!/bin/bash
mkdir -p gpurun_out
CMD="python benchmarks/conv_layers.py --ops fprop --only L0.stem,L0.c1 --reps 1"
$CMD > gpurun_out/ncu2_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tap_gemm" -s 6 -c 2 -o gpurun_out/prof_stem $CMD > gpurun_out/ncu2b.log 2>&1
ls -la gpurun_out/prof_stem.ncu-rep; tail -3 gpurun_out/ncu2b.log
