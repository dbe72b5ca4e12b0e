#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_conv.py -q -x > gpurun_out/r2d_conv_tests.log 2>&1; echo "exit $?" >> gpurun_out/r2d_conv_tests.log
tail -15 gpurun_out/r2d_conv_tests.log | cut -c1-300
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; }
python benchmarks/conv_layers.py --reps 5 --only L0.c1,L0.dec0,L1.c0 --ops wgrad > gpurun_out/r2d_pair.log 2>&1; echo pair; show gpurun_out/r2d_pair.log
B200SEG_WGRAD_PAIR=0 python benchmarks/conv_layers.py --reps 5 --only L0.c1,L0.dec0,L1.c0 --ops wgrad > gpurun_out/r2d_nopair.log 2>&1; echo PAIR=0; show gpurun_out/r2d_nopair.log
