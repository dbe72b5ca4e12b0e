#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x > gpurun_out/r2g_conv_tests.log 2>&1; tail -2 gpurun_out/r2g_conv_tests.log
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; }
L="L2.dec0,L3.dec0,L4.c0,L4.dec0,L5.c0"
python benchmarks/conv_layers.py --reps 10 --only $L --ops wgrad > gpurun_out/r2g_def.log 2>&1; echo default; show gpurun_out/r2g_def.log
B200SEG_WGRAD_WAVES=1 python benchmarks/conv_layers.py --reps 10 --only $L --ops wgrad > gpurun_out/r2g_w1.log 2>&1; echo WAVES=1; show gpurun_out/r2g_w1.log
B200SEG_WGRAD_VEC=0 python benchmarks/conv_layers.py --reps 10 --only $L --ops wgrad > gpurun_out/r2g_v0.log 2>&1; echo VEC=0; show gpurun_out/r2g_v0.log
B200SEG_WGRAD_MT=1 python benchmarks/conv_layers.py --reps 10 --only $L --ops wgrad > gpurun_out/r2g_mt1.log 2>&1; echo MT=1; show gpurun_out/r2g_mt1.log
