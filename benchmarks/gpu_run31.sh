#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_elem.py tests/test_gpu_train.py -q -x > gpurun_out/tests31.log 2>&1; echo "exit $?" >> gpurun_out/tests31.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench31.json 2> gpurun_out/bench31.err
timeout 600 python benchmarks/profile_step.py --impl own --out gpurun_out/step31.txt > gpurun_out/step31.log 2>&1
tail -5 gpurun_out/tests31.log; python - <<'PY'
import json
for f in ('bench31',):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'))
PY
grep -E "bn_|total|channel_stats" gpurun_out/step31.txt
