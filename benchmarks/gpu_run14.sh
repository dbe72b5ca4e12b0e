#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -s -k allreduce > gpurun_out/tests14.log 2>&1; echo "exit $?" >> gpurun_out/tests14.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 benchmarks/allreduce_sweep.py --max-mb 512 --out gpurun_out/allreduce_sweep_2gpu.jsonl > gpurun_out/sweep14.log 2>&1
tail -6 gpurun_out/tests14.log; grep bytes gpurun_out/sweep14.log | cut -c1-600
