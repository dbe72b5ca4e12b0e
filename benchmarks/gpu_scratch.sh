#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t9_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/t9_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/t9_smoke.log 2>&1; echo "build+smoke exit $?"; grep "smoke ok" gpurun_out/t9_smoke.log | cut -c1-200
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/t9_bench.json 2> gpurun_out/t9_bench.err; cut -c1-170 gpurun_out/t9_bench.json; grep -o '"e2e".*' gpurun_out/t9_bench.json | cut -c1-120
