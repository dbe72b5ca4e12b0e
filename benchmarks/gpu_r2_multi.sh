#!/bin/bash
# N GPUs ($1): multi-GPU tests, then bench own (overlapped exchange, default) and own --no-overlap
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_multi.py -q -x -s > gpurun_out/r2_multi_tests_$N.log 2>&1; echo "exit $?" >> gpurun_out/r2_multi_tests_$N.log
grep -E "passed|failed|exit|Error|error|assert" gpurun_out/r2_multi_tests_$N.log | cut -c1-300 | tail -15
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 10 --warmup 3 "${@:3}" > gpurun_out/r2_multi_bench_${N}_$2.json 2> gpurun_out/r2_multi_bench_${N}_$2.err; }
run 29521 ovl
run 29522 noovl --no-overlap
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_multi_bench_${N}_single.json 2> gpurun_out/r2_multi_bench_${N}_single.err
for f in ovl noovl single; do python -c "
import json
try:
    d=json.loads(open('gpurun_out/r2_multi_bench_${N}_$f.json').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'),d.get('e2e',{}).get('ms_per_step'),d['config'].get('allreduce'),d['config'].get('grad_buckets'),d.get('clocks'))
except Exception as e: print('$f failed',e)"; done
tail -3 gpurun_out/r2_multi_bench_${N}_ovl.err
