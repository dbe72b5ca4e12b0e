#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/r2_tests_1gpu.log 2>&1; echo "exit $?" >> gpurun_out/r2_tests_1gpu.log
tail -4 gpurun_out/r2_tests_1gpu.log
L="L0.stem,L0.c1,L0.dec0"
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; }
python benchmarks/conv_layers.py --reps 5 --only $L --ops fprop,dgrad > gpurun_out/r2c_default.log 2>&1; echo default; show gpurun_out/r2c_default.log
B200SEG_ROWS4=2 python benchmarks/conv_layers.py --reps 5 --only $L --ops fprop > gpurun_out/r2c_rows4.log 2>&1; echo ROWS4=2; show gpurun_out/r2c_rows4.log
B200SEG_EPI_PRIVATE=0 python benchmarks/conv_layers.py --reps 5 --only $L --ops fprop,dgrad > gpurun_out/r2c_shared.log 2>&1; echo EPI_PRIVATE=0; show gpurun_out/r2c_shared.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/r2c_bench1.json 2> gpurun_out/r2c_bench1.err; cut -c1-200 gpurun_out/r2c_bench1.json; tail -3 gpurun_out/r2c_bench1.err
