#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -q -x > gpurun_out/tests38.log 2>&1; echo "exit $?" >> gpurun_out/tests38.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench38.json 2> gpurun_out/bench38.err
timeout 600 python bench.py --steps 10 --warmup 3 --impl reference > gpurun_out/bench38_ref.json 2> gpurun_out/bench38_ref.err
tail -3 gpurun_out/tests38.log; python - <<'PY'
import json
for f in ('bench38','bench38_ref'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'),d.get('e2e'))
PY
tail -3 gpurun_out/bench38.err
