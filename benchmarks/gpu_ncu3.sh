#!/bin/bash
# ncu evidence: (1) every launch of ONE training step, (2) --set full on the hot kernels of that step.
mkdir -p gpurun_out
CMD="python benchmarks/ncu_target.py --steps 3"
$CMD > gpurun_out/ncu3_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu3_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches3.csv $CMD > gpurun_out/ncu3_a.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"tap_gemm|wgrad_strip|wgrad_kernel" -c 75 -o gpurun_out/prof3_conv $CMD > gpurun_out/ncu3_b.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"bn_bwd_apply|bn_bwd_reduce|bn_relu_pool_fwd|head_ce|fused_adam|channel_stats|prep_input" -c 60 -o gpurun_out/prof3_elem $CMD > gpurun_out/ncu3_c.log 2>&1
ls -la gpurun_out/prof3_*.ncu-rep; wc -l gpurun_out/launches3.csv; tail -2 gpurun_out/ncu3_a.log gpurun_out/ncu3_b.log gpurun_out/ncu3_c.log
