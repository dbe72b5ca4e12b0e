#!/bin/bash
mkdir -p gpurun_out
B200SEG_PDL=1 timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_elem.py tests/test_gpu_train.py -q -x > gpurun_out/tests34.log 2>&1; echo "exit $?" >> gpurun_out/tests34.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench34_off.json 2> gpurun_out/bench34_off.err
B200SEG_PDL=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench34_on.json 2> gpurun_out/bench34_on.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench34_off2.json 2> gpurun_out/bench34_off2.err
B200SEG_PDL=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench34_on2.json 2> gpurun_out/bench34_on2.err
tail -4 gpurun_out/tests34.log
for f in off on off2 on2; do python -c "
import json
d=json.loads(open('gpurun_out/bench34_$f.json').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'))"; done; tail -3 gpurun_out/bench34_on.err
