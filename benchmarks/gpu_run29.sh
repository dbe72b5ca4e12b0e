#!/bin/bash
mkdir -p gpurun_out
timeout 600 python benchmarks/profile_step.py --impl own --out gpurun_out/step29.txt > gpurun_out/step29.log 2>&1
head -32 gpurun_out/step29.txt
