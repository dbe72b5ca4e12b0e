#!/bin/bash
# A/B of the overlapped exchange at N GPUs: CTAs of the bucket reduce kernels, and exchange-after-backward
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -q -x -k stalled > gpurun_out/r2_peerloss_$N.log 2>&1; tail -2 gpurun_out/r2_peerloss_$N.log
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 30 --warmup 5 --no-e2e "$@" > gpurun_out/r2_ab_${N}_$tag.json 2> gpurun_out/r2_ab_${N}_$tag.err
python -c "
import json
try:
    d=json.loads(open('gpurun_out/r2_ab_${N}_$tag.json').read().strip().split('\n')[-1]);print('$tag',d.get('value'),d.get('ms_per_step'),d['config'].get('allreduce'),d.get('clocks'))
except Exception as e: print('$tag failed',e)"; }
B200SEG_OVL_BLOCKS=16 run ovl16
run noovl --no-overlap
B200SEG_OVL_BLOCKS=4 run ovl4
B200SEG_OVL_BLOCKS=2 run ovl2
run noovl_b --no-overlap
B200SEG_OVL_BLOCKS=8 run ovl8
