#!/bin/bash
# 2-GPU sanity of the final build: own arm bench (exchange under backward + tail kernel + pipelined readback) and the overlap test
N=2
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/t6_bench_2_own.json 2> gpurun_out/t6_bench_2_own.err; echo "bench exit $?"; cut -c1-260 gpurun_out/t6_bench_2_own.json; grep -o '"e2e".*' gpurun_out/t6_bench_2_own.json | cut -c1-200; tail -2 gpurun_out/t6_bench_2_own.err
timeout 300 python -m pytest tests/test_gpu_multi.py -q -x -k "overlapped" > gpurun_out/t6_multi_2.log 2>&1; echo "multi exit $?"; tail -2 gpurun_out/t6_multi_2.log
