#!/bin/bash
N=4
mkdir -p gpurun_out
for U in 4 8; do
B200SEG_NVLS_UNROLL=$U timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/allreduce_sweep.py --min-mb 64 --max-mb 256 --nvls-only --grid-sweep --blocks 8,16,32,64 --out gpurun_out/s_nvls_u$U.jsonl > gpurun_out/s_nvls_u$U.log 2>&1
python - <<PY
import json
for l in open('gpurun_out/s_nvls_u$U.jsonl'):
    r=json.loads(l)
    print('U=$U', r['bytes']>>20,'MB', {k[:-7]:v for k,v in r.items() if k.endswith('_busGBs')})
PY
done
