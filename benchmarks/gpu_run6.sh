#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x > gpurun_out/tests6.log 2>&1; echo "exit $?" >> gpurun_out/tests6.log
timeout 900 python benchmarks/conv_layers.py --out gpurun_out/conv_layers6.jsonl > gpurun_out/conv_layers6.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench6.json 2> gpurun_out/bench6.err
tail -3 gpurun_out/tests6.log; cut -c1-175 gpurun_out/conv_layers6.log | tail -48; cat gpurun_out/bench6.json; tail -3 gpurun_out/bench6.err
