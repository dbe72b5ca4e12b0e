#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_elem.py tests/test_gpu_train.py -q -x > gpurun_out/tests20.log 2>&1; echo "exit $?" >> gpurun_out/tests20.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench20.json 2> gpurun_out/bench20.err
# other BASELINE configs (1 GPU): reference width 32; 2x width 128; 1024x1024 small batch
for cfg in "32 512 16" "128 512 16" "64 1024 2" "32 512 1"; do set -- $cfg
  timeout 600 python bench.py --steps 5 --warmup 3 --base-channels $1 --tile $2 --batch $3 > gpurun_out/bench20_own_b$1_t$2_n$3.json 2> gpurun_out/bench20_own_b$1_t$2_n$3.err
  timeout 600 python bench.py --impl reference --steps 5 --warmup 3 --base-channels $1 --tile $2 --batch $3 > gpurun_out/bench20_ref_b$1_t$2_n$3.json 2> gpurun_out/bench20_ref_b$1_t$2_n$3.err
done
tail -3 gpurun_out/tests20.log
for f in gpurun_out/bench20*.json; do python -c "
import json,sys
try:
  d=json.loads(open('$f').read().strip().split('\n')[-1]);print('$f',d.get('impl'),d.get('value'),d.get('ms_per_step'),(d.get('e2e') or {}).get('value'))
except Exception as e: print('$f','ERR',e)"; done
