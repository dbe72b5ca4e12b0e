#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_gpu_train.py -q -s -k "headline or parity" > gpurun_out/r2j_tests.log 2>&1; echo "exit $?" >> gpurun_out/r2j_tests.log
grep -E "passed|failed|^exit|last-20|worst|Error|assert" gpurun_out/r2j_tests.log | cut -c1-400 | tail -12
tail -5 gpurun_out/convergence_parity.txt 2>/dev/null
sort -r gpurun_out/grad_errors_headline_shape.txt 2>/dev/null | head -8
