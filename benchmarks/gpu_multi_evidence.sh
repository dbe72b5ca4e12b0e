#!/bin/bash
# N-GPU lease ($1): multi-GPU tests (raw log), both arms of the headline bench with e2e, all-reduce sweep
N=${1:-8}
mkdir -p gpurun_out
if [ "$2" != "notests" ]; then
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x -s > gpurun_out/fin_multigpu_pytest_$N.log 2>&1; echo "exit $?" >> gpurun_out/fin_multigpu_pytest_$N.log
grep -E "passed|failed|^exit" gpurun_out/fin_multigpu_pytest_$N.log | tail -3
fi
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/fin_bench_${N}_$tag.json 2> gpurun_out/fin_bench_${N}_$tag.err
python -c "
import json
try:
    d=json.loads(open('gpurun_out/fin_bench_${N}_$tag.json').read().strip().split('\n')[-1]);print('$tag',d.get('value'),d.get('ms_per_step'),(d.get('e2e') or {}).get('value'),d['config'].get('allreduce'),d.get('clocks'))
except Exception as e: print('$tag failed',e)"; }
run ref --impl reference
run own
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/allreduce_sweep.py --max-mb 256 --grid-sweep --out gpurun_out/fin_sweep_$N.jsonl > gpurun_out/fin_sweep_$N.log 2>&1
python - <<PY
import json
for l in open('gpurun_out/fin_sweep_$N.jsonl'):
    r=json.loads(l)
    print(r['bytes']>>10,'KB', {k[:-7]:v for k,v in r.items() if k.endswith('_busGBs')})
PY
