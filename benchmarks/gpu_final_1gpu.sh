#!/bin/bash
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -q -x -m gpu > gpurun_out/fin_tests_1gpu.log 2>&1; echo "exit $?" >> gpurun_out/fin_tests_1gpu.log
tail -3 gpurun_out/fin_tests_1gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/fin_smoke.log 2>&1; tail -1 gpurun_out/fin_smoke.log | cut -c1-200
python benchmarks/conv_layers.py --reps 5 --out gpurun_out/fin_conv_layers_b64.jsonl > gpurun_out/fin_conv_layers_b64.log 2>&1; tail -1 gpurun_out/fin_conv_layers_b64.log
python benchmarks/conv_layers.py --base 128 --reps 3 --out gpurun_out/fin_conv_layers_b128.jsonl > gpurun_out/fin_conv_layers_b128.log 2>&1; tail -1 gpurun_out/fin_conv_layers_b128.log
python benchmarks/profile_step.py --impl own --graph --steps 10 --out gpurun_out/fin_step_own.txt > /dev/null 2>&1; head -1 gpurun_out/fin_step_own.txt
python benchmarks/profile_step.py --impl reference --steps 5 --out gpurun_out/fin_step_ref.txt > /dev/null 2>&1; head -1 gpurun_out/fin_step_ref.txt
timeout 600 python bench.py --impl reference > gpurun_out/fin_bench_1_ref.json 2> gpurun_out/fin_bench_1_ref.err; cut -c1-150 gpurun_out/fin_bench_1_ref.json
timeout 600 python bench.py > gpurun_out/fin_bench_1_own.json 2> gpurun_out/fin_bench_1_own.err; cut -c1-150 gpurun_out/fin_bench_1_own.json
timeout 200 python benchmarks/tail_bench.py > gpurun_out/fin_tail_bench.json 2> gpurun_out/fin_tail_bench.err; cat gpurun_out/fin_tail_bench.json
B200SEG_HEAD_TAIL=0 timeout 600 python bench.py > gpurun_out/fin_bench_1_own_notail.json 2> gpurun_out/fin_bench_1_own_notail.err; cut -c1-150 gpurun_out/fin_bench_1_own_notail.json
