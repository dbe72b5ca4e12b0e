#!/bin/bash
# every BASELINE.json config on 1 GPU, both arms  -> gpurun_out/table_*.json
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python bench.py --steps 10 --warmup 3 "$@" > gpurun_out/table_${name}_own.json 2> gpurun_out/table_${name}_own.err; timeout 900 python bench.py --steps 10 --warmup 3 --impl reference "$@" > gpurun_out/table_${name}_ref.json 2> gpurun_out/table_${name}_ref.err; }
run b64
run b128 --base-channels 128
run t1024 --tile 1024 --batch 2
run b32 --base-channels 32
run b32n1 --base-channels 32 --batch 1
python - <<'PY'
import json
for n in ('b64','b128','t1024','b32','b32n1'):
    r=[]
    for arm in ('ref','own'):
        try:
            d=json.loads(open(f'gpurun_out/table_{n}_{arm}.json').read().strip().split('\n')[-1])
            r.append(f"{arm} {d['value']:.1f} ({d['ms_per_step']:.2f}) e2e {d.get('e2e',{}).get('value')}")
        except Exception as e:
            r.append(f"{arm} ERR {e}")
    print(n, ' | '.join(r))
PY
