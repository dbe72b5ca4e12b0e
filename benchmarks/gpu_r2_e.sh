#!/bin/bash
mkdir -p gpurun_out
python benchmarks/profile_step.py --impl own --out gpurun_out/r2e_step_kernels.txt > gpurun_out/r2e_prof.log 2>&1; head -32 gpurun_out/r2e_step_kernels.txt | cut -c1-150
python benchmarks/conv_layers.py --reps 5 --out gpurun_out/r2e_conv_layers.jsonl > gpurun_out/r2e_conv_layers.log 2>&1; tail -1 gpurun_out/r2e_conv_layers.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2e_bench1.json 2> gpurun_out/r2e_bench1.err; cut -c1-160 gpurun_out/r2e_bench1.json
