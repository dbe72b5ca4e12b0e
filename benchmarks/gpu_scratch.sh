#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t3_tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/t3_tests.log
timeout 200 python benchmarks/tail_bench.py > gpurun_out/t3_tail_bench.json 2> gpurun_out/t3_tail_bench.err; cat gpurun_out/t3_tail_bench.json; tail -3 gpurun_out/t3_tail_bench.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t3_smoke_import.log 2>&1; echo "smoke(import) exit $?"; grep -v "^frame" gpurun_out/t3_smoke_import.log | tail -5 | cut -c1-400
timeout 600 python __graft_entry__.py smoke > gpurun_out/t3_smoke_main.log 2>&1; echo "smoke(main) exit $?"; grep -v "^frame" gpurun_out/t3_smoke_main.log | tail -12 | cut -c1-400
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/t3_bench_tail.json 2> gpurun_out/t3_bench_tail.err; cut -c1-200 gpurun_out/t3_bench_tail.json; tail -3 gpurun_out/t3_bench_tail.err
