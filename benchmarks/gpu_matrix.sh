#!/bin/bash
# BASELINE.json configs 3, 4 and the reference-as-shipped config at N GPUs ($1), both arms
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; if [ $N -gt 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551"; else L=python; fi
timeout 600 $L bench.py --gpus $N --steps 10 --warmup 3 "$@" > gpurun_out/mx_${N}_$tag.json 2> gpurun_out/mx_${N}_$tag.err
python -c "
import json
try:
    d=json.loads(open('gpurun_out/mx_${N}_$tag.json').read().strip().split('\n')[-1]);print('$tag',d.get('value'),d.get('ms_per_step'),(d.get('e2e') or {}).get('value'),d.get('clocks',{}).get('reasons'))
except Exception as e: print('$tag failed',e)"; }
for P in ${2:-b128 tile1024 shipped}; do
run ${P}_ref --preset $P --impl reference
run ${P}_own --preset $P
done
