#!/bin/bash
# N GPUs (passed as $1): multi-GPU tests, then bench own (native), own (fused), reference
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x -s > gpurun_out/multi_tests_$N.log 2>&1; echo "exit $?" >> gpurun_out/multi_tests_$N.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 10 --warmup 3 "${@:3}" > gpurun_out/multi_bench_${N}_$2.json 2> gpurun_out/multi_bench_${N}_$2.err; }
run 29521 own
# (fused arm: see profiles/bench_8gpu.txt; skipped here to save box time)
run 29523 ref --impl reference
grep -E "passed|failed|exit|errs" gpurun_out/multi_tests_$N.log | cut -c1-300
for f in own ref; do python -c "
import json
d=json.loads(open('gpurun_out/multi_bench_${N}_$f.json').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'),d.get('e2e'),d.get('clocks'))"; done
tail -2 gpurun_out/multi_bench_${N}_own.err
