#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k wgrad > gpurun_out/tests22.log 2>&1; echo "exit $?" >> gpurun_out/tests22.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench22.json 2> gpurun_out/bench22.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --base-channels 128 > gpurun_out/bench22_b128.json 2> gpurun_out/bench22_b128.err
tail -3 gpurun_out/tests22.log
for f in gpurun_out/bench22.json gpurun_out/bench22_b128.json; do python -c "
import json
d=json.loads(open('$f').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'))"; done
