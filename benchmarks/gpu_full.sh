#!/bin/bash
# full single-GPU check: every GPU test, smoke(), bench both arms
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/full_tests.log 2>&1; echo "exit $?" >> gpurun_out/full_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/full_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/full_bench_own.json 2> gpurun_out/full_bench_own.err
timeout 600 python bench.py --impl reference > gpurun_out/full_bench_ref.json 2> gpurun_out/full_bench_ref.err
tail -4 gpurun_out/full_tests.log; tail -2 gpurun_out/full_smoke.log; cat gpurun_out/full_bench_own.json; cat gpurun_out/full_bench_ref.json
