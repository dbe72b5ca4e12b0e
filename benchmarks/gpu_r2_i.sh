#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -q -x -m gpu -s > gpurun_out/r2i_tests_1gpu.log 2>&1; echo "exit $?" >> gpurun_out/r2i_tests_1gpu.log
grep -E "passed|failed|^exit|last-20|largest relative|Error|assert" gpurun_out/r2i_tests_1gpu.log | cut -c1-400 | tail -12
cat gpurun_out/convergence_parity.txt 2>/dev/null | tail -8
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench1.json 2> gpurun_out/r2i_bench1.err; cut -c1-160 gpurun_out/r2i_bench1.json
