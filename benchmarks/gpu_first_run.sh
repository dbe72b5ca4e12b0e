#!/bin/bash
# First GPU bring-up: descriptor probe, kernel tests, first bench numbers.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
P=build/umma_probe
if [ ! -x $P ]; then mkdir -p build; nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -Icsrc -o $P csrc/standalone_umma_probe.cu; fi
: > gpurun_out/probe.log
run() { timeout 60 $P "$@" >> gpurun_out/probe.log 2>&1 || echo "  (exit $? for: $*)" >> gpurun_out/probe.log; }
for n in 64 128 256; do run 0 $n; done
for n in 64 128 256; do run 1 $n; done
run 1 128 0 1024 8192
for n in 64 128; do run 2 $n; done
run 2 128 0 1024 8192
for s in 1 2 3 5; do run 3 128 $s; run 4 128 $s; done
for s in 1 2 3 5; do run 5 128 $s; run 6 128 $s; done
timeout 900 python -m pytest tests/test_gpu_elem.py -q > gpurun_out/elem.log 2>&1
echo "elem exit $?" >> gpurun_out/elem.log
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k fprop > gpurun_out/conv_fprop.log 2>&1
echo "exit $?" >> gpurun_out/conv_fprop.log
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k dgrad > gpurun_out/conv_dgrad.log 2>&1
echo "exit $?" >> gpurun_out/conv_dgrad.log
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k wgrad > gpurun_out/conv_wgrad.log 2>&1
echo "exit $?" >> gpurun_out/conv_wgrad.log
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k convT > gpurun_out/conv_convT.log 2>&1
echo "exit $?" >> gpurun_out/conv_convT.log
timeout 900 python -m pytest tests/test_gpu_train.py -q -k "cudnn or trainer" > gpurun_out/train.log 2>&1
echo "exit $?" >> gpurun_out/train.log
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 python bench.py --conv cudnn --steps 5 --warmup 3 > gpurun_out/bench_own_cudnn.json 2> gpurun_out/bench_own_cudnn.err
tail -3 gpurun_out/*.log; cat gpurun_out/bench_*.json
