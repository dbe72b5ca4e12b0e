#!/bin/bash
mkdir -p gpurun_out
for i in 1 2 3; do timeout 600 python -m pytest tests/test_gpu_train.py -q -x -k headline -s > gpurun_out/s_t$i.log 2>&1; grep -E "passed|failed|worst|Assertion" gpurun_out/s_t$i.log | cut -c1-250 | tail -3; done
sort -r gpurun_out/grad_errors_headline_shape.txt | head -5
