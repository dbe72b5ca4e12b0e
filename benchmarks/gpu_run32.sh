#!/bin/bash
# 2 GPUs: fused wgrad+all-reduce test, then bench native vs fused
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x -s > gpurun_out/tests32.log 2>&1; echo "exit $?" >> gpurun_out/tests32.log
for comm in native fused; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --comm $comm > gpurun_out/bench32_$comm.json 2> gpurun_out/bench32_$comm.err
done
tail -15 gpurun_out/tests32.log
for comm in native fused; do python -c "
import json
d=json.loads(open('gpurun_out/bench32_$comm.json').read().strip().split('\n')[-1]);print('$comm',d.get('value'),d.get('ms_per_step'),d['config'].get('comm'))"; tail -2 gpurun_out/bench32_$comm.err; done
