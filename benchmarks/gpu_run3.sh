#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py -q > gpurun_out/train3.log 2>&1; echo "exit $?" >> gpurun_out/train3.log
timeout 600 python benchmarks/profile_step.py --impl own --conv cudnn --out gpurun_out/prof_own_cudnn.txt > /dev/null 2> gpurun_out/prof1.err
timeout 600 python benchmarks/profile_step.py --impl own --conv tcgen05 --out gpurun_out/prof_own_tc.txt > /dev/null 2> gpurun_out/prof2.err
timeout 600 python benchmarks/profile_step.py --impl reference --out gpurun_out/prof_ref.txt > /dev/null 2> gpurun_out/prof3.err
tail -4 gpurun_out/train3.log; cat gpurun_out/prof_own_cudnn.txt; cat gpurun_out/prof_own_tc.txt | head -30; cat gpurun_out/prof_ref.txt | head -45
