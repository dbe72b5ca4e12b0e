#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/tests19.log 2>&1; echo "exit $?" >> gpurun_out/tests19.log
timeout 600 python benchmarks/profile_step.py --impl own --conv tcgen05 --out gpurun_out/prof19.txt > /dev/null 2> gpurun_out/prof19.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench19.json 2> gpurun_out/bench19.err
tail -3 gpurun_out/tests19.log; cut -c1-150 gpurun_out/prof19.txt | head -20; cat gpurun_out/bench19.json
