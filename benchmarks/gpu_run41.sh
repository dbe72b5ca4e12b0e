#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests41.log 2>&1; echo "exit $?" >> gpurun_out/tests41.log
for i in 1 2; do
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench41_b64_$i.json 2> gpurun_out/bench41.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e --base-channels 128 > gpurun_out/bench41_b128_$i.json 2> gpurun_out/bench41b.err
done
tail -3 gpurun_out/tests41.log
for f in b64_1 b128_1 b64_2 b128_2; do python -c "
import json
d=json.loads(open('gpurun_out/bench41_$f.json').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'))"; done
