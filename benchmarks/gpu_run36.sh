#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests36.log 2>&1; echo "exit $?" >> gpurun_out/tests36.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench36.json 2> gpurun_out/bench36.err
tail -6 gpurun_out/tests36.log; python - <<'PY'
import json
for f in ('bench36',):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'),d.get('gpu_launches'))
PY
tail -3 gpurun_out/bench36.err
