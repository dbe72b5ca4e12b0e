#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_multi.py -q -x -s > gpurun_out/r2k_multi_tests_$N.log 2>&1; echo "exit $?" >> gpurun_out/r2k_multi_tests_$N.log
grep -E "passed|failed|^exit|Error|assert" gpurun_out/r2k_multi_tests_$N.log | cut -c1-300 | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 benchmarks/allreduce_sweep.py --max-mb 256 --grid-sweep --out gpurun_out/r2k_sweep_$N.jsonl > gpurun_out/r2k_sweep_$N.log 2>&1
python - <<PY
import json
for l in open('gpurun_out/r2k_sweep_$N.jsonl'):
    r=json.loads(l)
    print(r['bytes']>>10,'KB', {k[:-7]:v for k,v in r.items() if k.endswith('_busGBs')})
PY
