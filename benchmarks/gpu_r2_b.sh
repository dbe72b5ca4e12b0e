#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/r2_tests_1gpu.log 2>&1; echo "exit $?" >> gpurun_out/r2_tests_1gpu.log
tail -4 gpurun_out/r2_tests_1gpu.log
python benchmarks/conv_layers.py --reps 5 --only L0.stem,L0.c1,L0.dec0,L1.c0 > gpurun_out/r2_l0_default.log 2>&1; grep -v totals gpurun_out/r2_l0_default.log | cut -c1-150
B200SEG_ROWS4=2 python benchmarks/conv_layers.py --reps 5 --only L0.stem,L0.c1 --ops fprop > gpurun_out/r2_l0_rows4.log 2>&1; echo ROWS4=2; grep -v totals gpurun_out/r2_l0_rows4.log | cut -c1-150
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; cat gpurun_out/r2_bench1.json; tail -3 gpurun_out/r2_bench1.err
