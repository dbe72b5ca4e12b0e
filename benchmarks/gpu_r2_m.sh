#!/bin/bash
mkdir -p gpurun_out
show() { grep -v totals $1 | python -c "
import sys,json
for l in sys.stdin:
    try: r=json.loads(l)
    except Exception: continue
    print(f\"  {r['layer']:8s} {r['op']:6s} {r['HW']:4d} {r['cin']:5d} {r['cout']:5d} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f}\")"; tail -1 $1; }
python benchmarks/conv_layers.py --base 128 --reps 3 > gpurun_out/r2m_b128_def.log 2>&1; echo default; show gpurun_out/r2m_b128_def.log
B200SEG_WGRAD_PAIR=0 python benchmarks/conv_layers.py --base 128 --reps 3 --ops wgrad > gpurun_out/r2m_b128_nopair.log 2>&1; echo PAIR=0; show gpurun_out/r2m_b128_nopair.log
