#!/bin/bash
# 8-GPU lease: multi-GPU tests (raw log kept), overlap A/B, both arms of the headline bench
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x -s -m "gpu" > gpurun_out/r2_multigpu_pytest_$N.log 2>&1; echo "exit $?" >> gpurun_out/r2_multigpu_pytest_$N.log
grep -E "passed|failed|^exit" gpurun_out/r2_multigpu_pytest_$N.log | tail -3
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 30 --warmup 5 "$@" > gpurun_out/r2_ab_${N}_$tag.json 2> gpurun_out/r2_ab_${N}_$tag.err
python -c "
import json
try:
    d=json.loads(open('gpurun_out/r2_ab_${N}_$tag.json').read().strip().split('\n')[-1]);print('$tag',d.get('value'),d.get('ms_per_step'),(d.get('e2e') or {}).get('ms_per_step'),d['config'].get('allreduce'),d.get('clocks'))
except Exception as e: print('$tag failed',e)"; }
B200SEG_OVL_BLOCKS=16 run ovl16 --no-e2e
run noovl --no-overlap --no-e2e
B200SEG_OVL_BLOCKS=4 run ovl4 --no-e2e
run noovl_b --no-overlap --no-e2e
timeout 300 python bench.py --steps 30 --warmup 5 --no-e2e > gpurun_out/r2_ab_${N}_single.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_ab_${N}_single.json').read().strip().split('\n')[-1]); print('single', d['value'], d['ms_per_step'], d['clocks'])"
