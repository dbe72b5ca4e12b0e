#!/bin/bash
# round 2, call 1: source-level ncu of the three kernels the verdict singles out (per-layer harness, one launch each)
mkdir -p gpurun_out
cap() {   # name, kernel regex, layer, op
  CMD="python benchmarks/conv_layers.py --only $3 --ops $4 --reps 1"
  $CMD > gpurun_out/r2_$1_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$2 -s 3 -c 1 -o gpurun_out/r2_$1 $CMD > gpurun_out/r2_$1_ncu.log 2>&1
  tail -2 gpurun_out/r2_$1_plain.log; tail -1 gpurun_out/r2_$1_ncu.log
}
cap fprop_l0c1 tap_gemm L0.c1 fprop
cap wgrad_l4c0 wgrad_kernel L4.c0 wgrad
cap wgrad_l0c1 wgrad_strip L0.c1 wgrad
cap fprop_l0dec0 tap_gemm L0.dec0 fprop
python benchmarks/conv_layers.py --reps 5 --out gpurun_out/r2_conv_layers_base.jsonl > gpurun_out/r2_conv_layers_base.log 2>&1
tail -3 gpurun_out/r2_conv_layers_base.log
ls -la gpurun_out/r2_*.ncu-rep
