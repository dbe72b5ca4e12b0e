#!/bin/bash
# 2 GPUs: per-kernel profile, native vs fused
mkdir -p gpurun_out
for comm in native fused; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 benchmarks/profile_step.py --comm $comm --out gpurun_out/step33_$comm.txt > gpurun_out/step33_$comm.log 2>&1
done
for comm in native fused; do echo "== $comm"; head -14 gpurun_out/step33_$comm.txt | cut -c1-100; grep -E "allreduce|barrier" gpurun_out/step33_$comm.txt | cut -c1-100; done
tail -3 gpurun_out/step33_fused.log
