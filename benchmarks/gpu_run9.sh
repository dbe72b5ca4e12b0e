#!/bin/bash
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/tests9.log 2>&1; echo "exit $?" >> gpurun_out/tests9.log
for comm in torch native; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --comm $comm > gpurun_out/bench9_$comm.json 2> gpurun_out/bench9_$comm.err
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --impl reference > gpurun_out/bench9_ref.json 2> gpurun_out/bench9_ref.err
tail -15 gpurun_out/tests9.log; cat gpurun_out/bench9_*.json; tail -5 gpurun_out/bench9_native.err
