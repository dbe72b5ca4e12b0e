#!/bin/bash
# 8-GPU validation: all-reduce tests, sweep vs NCCL, scaling bench (own native / own NCCL / reference)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -q -x -s -k allreduce > gpurun_out/tests16.log 2>&1; echo "exit $?" >> gpurun_out/tests16.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 benchmarks/allreduce_sweep.py --max-mb 512 --out gpurun_out/allreduce_sweep_8gpu.jsonl > gpurun_out/sweep16.log 2>&1
for comm in native torch; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 10 --warmup 3 --comm $comm --no-e2e > gpurun_out/bench16_8_$comm.json 2> gpurun_out/bench16_8_$comm.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --steps 10 --warmup 3 --impl reference --no-e2e > gpurun_out/bench16_8_ref.json 2> gpurun_out/bench16_8_ref.err
tail -4 gpurun_out/tests16.log; grep bytes gpurun_out/sweep16.log | cut -c1-700; cat gpurun_out/bench16_8_*.json | cut -c1-300; tail -3 gpurun_out/bench16_8_native.err
