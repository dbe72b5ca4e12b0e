#!/bin/bash
mkdir -p gpurun_out
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; }
L="L1.c1,L1.dec0,L2.c0,L2.c1,L2.dec0,L3.c0,L3.c1,L3.dec0"
python benchmarks/conv_layers.py --reps 5 --only $L --ops wgrad > gpurun_out/r2f_def.log 2>&1; echo default; show gpurun_out/r2f_def.log
B200SEG_WGRAD_PAIR=2 python benchmarks/conv_layers.py --reps 5 --only $L --ops wgrad > gpurun_out/r2f_pair.log 2>&1; echo PAIR=2; show gpurun_out/r2f_pair.log
