#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_elem.py tests/test_gpu_train.py -q > gpurun_out/tests4.log 2>&1; echo "exit $?" >> gpurun_out/tests4.log
timeout 900 python benchmarks/conv_layers.py --ops fprop --out gpurun_out/conv_layers4.jsonl > gpurun_out/conv_layers4.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_tc4.json 2> gpurun_out/bench_own_tc4.err
tail -8 gpurun_out/tests4.log; cut -c1-200 gpurun_out/conv_layers4.log | tail -20; cat gpurun_out/bench_own_tc4.json; cat gpurun_out/grad_cosines_tcgen05.txt | sort | head -12
