#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests24.log 2>&1; echo "exit $?" >> gpurun_out/tests24.log
timeout 600 python benchmarks/conv_layers.py --ops fprop --only L0.stem,L0.c1 --out gpurun_out/conv_layers24.jsonl > gpurun_out/conv_layers24.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench24.json 2> gpurun_out/bench24.err
tail -5 gpurun_out/tests24.log; cut -c1-160 gpurun_out/conv_layers24.log | head -4
python -c "
import json
d=json.loads(open('gpurun_out/bench24.json').read().strip().split('\n')[-1]);print(d.get('value'),d.get('ms_per_step'))"; tail -3 gpurun_out/bench24.err
