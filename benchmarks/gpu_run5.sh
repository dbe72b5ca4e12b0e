#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x > gpurun_out/tests5.log 2>&1; echo "exit $?" >> gpurun_out/tests5.log
timeout 900 python benchmarks/conv_layers.py --ops wgrad --out gpurun_out/conv_layers5.jsonl > gpurun_out/conv_layers5.log 2>&1
timeout 600 python -m pytest tests/test_gpu_train.py -q > gpurun_out/train5.log 2>&1; echo "exit $?" >> gpurun_out/train5.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench5.json 2> gpurun_out/bench5.err
tail -8 gpurun_out/tests5.log; cut -c1-200 gpurun_out/conv_layers5.log | tail -20; tail -3 gpurun_out/train5.log; cat gpurun_out/bench5.json; tail -3 gpurun_out/bench5.err
