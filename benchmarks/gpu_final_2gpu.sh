#!/bin/bash
N=2
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/fin_bench_${N}_$tag.json 2> gpurun_out/fin_bench_${N}_$tag.err
python -c "
import json
try:
    d=json.loads(open('gpurun_out/fin_bench_${N}_$tag.json').read().strip().split('\n')[-1]);print('$tag',d.get('value'),d.get('ms_per_step'),(d.get('e2e') or {}).get('value'),d['config'].get('allreduce'),d.get('clocks'))
except Exception as e: print('$tag failed',e)"; }
run ref --impl reference
run own
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -k "overlapped or oneshot or stalled" > gpurun_out/fin_multi_2.log 2>&1; tail -2 gpurun_out/fin_multi_2.log
