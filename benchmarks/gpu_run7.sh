#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_elem.py tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests7.log 2>&1; echo "exit $?" >> gpurun_out/tests7.log
timeout 600 python benchmarks/profile_step.py --impl own --conv tcgen05 --out gpurun_out/prof7.txt > /dev/null 2> gpurun_out/prof7.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench7.json 2> gpurun_out/bench7.err
tail -3 gpurun_out/tests7.log; cut -c1-170 gpurun_out/prof7.txt | head -32; cat gpurun_out/bench7.json; tail -3 gpurun_out/bench7.err
