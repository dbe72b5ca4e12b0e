#!/bin/bash
mkdir -p gpurun_out
for r in 0 3 7 1; do
  B200SEG_REVERSE=$r python benchmarks/profile_step.py --impl own --graph --steps 10 --out gpurun_out/t5_step_rev$r.txt > /dev/null 2>&1; head -2 gpurun_out/t5_step_rev$r.txt | tail -1
done
