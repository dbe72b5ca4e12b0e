#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -s -k "overlapped" > gpurun_out/s_multi.log 2>&1; grep -E "passed|failed|comm_ms|Error|assert" gpurun_out/s_multi.log | cut -c1-600 | tail -6
