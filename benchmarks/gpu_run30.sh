#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_elem.py tests/test_gpu_train.py tests/test_zz_no_alloc.py -q -x > gpurun_out/tests30.log 2>&1; echo "exit $?" >> gpurun_out/tests30.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench30.json 2> gpurun_out/bench30.err
B200SEG_HEAD_FAST=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench30_slow.json 2> gpurun_out/bench30_b.err
timeout 600 python benchmarks/profile_step.py --impl own --out gpurun_out/step30.txt > gpurun_out/step30.log 2>&1
tail -5 gpurun_out/tests30.log; python - <<'PY'
import json
for f in ('bench30','bench30_slow'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'))
PY
grep -E "head_ce|total" gpurun_out/step30.txt
