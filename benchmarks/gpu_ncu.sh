#!/bin/bash
# ncu evidence: per-launch durations of one step, then --set full on the hot kernels.
mkdir -p gpurun_out
CMD="python benchmarks/ncu_target.py --steps 2"
$CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 440 -c 440 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tap_gemm|wgrad_strip|wgrad_kernel" -s 60 -c 12 -o gpurun_out/prof_conv $CMD > gpurun_out/ncu2.log 2>&1
$CMD > gpurun_out/ncu_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"bn_bwd_apply|bn_bwd_reduce|bn_relu_pool_fwd|head_ce|fused_adam" -s 30 -c 6 -o gpurun_out/prof_elem $CMD > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -3 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log
