#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/tests10.log 2>&1; echo "exit $?" >> gpurun_out/tests10.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --comm native > gpurun_out/bench10_native.json 2> gpurun_out/bench10_native.err
tail -15 gpurun_out/tests10.log; cat gpurun_out/bench10_*.json; tail -5 gpurun_out/bench10_native.err
