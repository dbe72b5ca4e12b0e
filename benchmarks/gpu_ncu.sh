#!/bin/bash
# ncu evidence, summarised ON the box (reports are too big to bring back whole):
#  (1) every launch of ONE training step, (2) --set full on the conv kernels of that step,
#  (3) --set full on the memory-bound kernels, (4) one small report with source for the top kernels.
mkdir -p gpurun_out /tmp/ncu
CMD="python benchmarks/ncu_target.py --steps 3"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_a.log 2>&1
ncu --profile-from-start off --set full --clock-control none -k regex:"tap_gemm|wgrad_strip|wgrad_kernel|wgrad_pair" -c 110 -o /tmp/ncu/conv $CMD > gpurun_out/ncu_b.log 2>&1
python benchmarks/ncu_summarize.py /tmp/ncu/conv.ncu-rep gpurun_out/ncu_conv.tsv "tcgen05 conv kernels, every launch of one training step (U-Net b=64, batch 16, 512x512)" >> gpurun_out/ncu_b.log 2>&1
ncu --profile-from-start off --set full --clock-control none -k regex:"bn_bwd_apply|bn_bwd_reduce|bn_relu_pool_fwd|head_tail|head_ce|fused_adam|channel_stats|prep_input" -c 36 -o /tmp/ncu/elem $CMD > gpurun_out/ncu_c.log 2>&1
python benchmarks/ncu_summarize.py /tmp/ncu/elem.ncu-rep gpurun_out/ncu_elem.tsv "memory-bound kernels of the same step (first 36 launches)" >> gpurun_out/ncu_c.log 2>&1
# small report with source: first strip fprop (two-row units), first strip dgrad, one wgrad_strip<128>, head
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"wgrad_pair|head_tail" -c 4 -o gpurun_out/prof_top $CMD > gpurun_out/ncu_d.log 2>&1
ncu -i gpurun_out/prof_top.ncu-rep --page details > gpurun_out/ncu_top_details.txt 2>&1
ls -la gpurun_out/prof_top.ncu-rep /tmp/ncu; wc -l gpurun_out/launches.csv gpurun_out/ncu_conv.tsv gpurun_out/ncu_elem.tsv
for f in a b c d; do tail -n 2 gpurun_out/ncu_$f.log; done
