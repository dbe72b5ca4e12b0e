#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k "wgrad or bench_shapes or alternative" > gpurun_out/r2l_conv_tests.log 2>&1; tail -3 gpurun_out/r2l_conv_tests.log | cut -c1-300
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; }
python benchmarks/conv_layers.py --reps 5 --ops wgrad > gpurun_out/r2l_wgrad.log 2>&1; show gpurun_out/r2l_wgrad.log; tail -1 gpurun_out/r2l_wgrad.log
