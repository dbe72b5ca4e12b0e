#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py -q > gpurun_out/train2.log 2>&1; echo "exit $?" >> gpurun_out/train2.log
timeout 900 python benchmarks/conv_layers.py --out gpurun_out/conv_layers.jsonl > gpurun_out/conv_layers.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_tc.json 2> gpurun_out/bench_own_tc.err
cat gpurun_out/conv_layers.log | tail -50; cat gpurun_out/bench_own_tc.json; tail -5 gpurun_out/train2.log
