#!/bin/bash
mkdir -p gpurun_out
timeout 900 python benchmarks/conv_layers.py --base 128 --reps 3 --out gpurun_out/conv_layers_b128.jsonl > gpurun_out/conv_layers_b128.log 2>&1
B200SEG_WGRAD_KERNEL=splitk timeout 900 python benchmarks/conv_layers.py --base 128 --reps 3 --ops wgrad --only L0.c1,L0.dec0,L1.c1 --out gpurun_out/conv_layers_b128_splitk.jsonl > /dev/null 2>&1
python - <<'PY'
import json
for f in ('gpurun_out/conv_layers_b128.jsonl','gpurun_out/conv_layers_b128_splitk.jsonl'):
    rows=[json.loads(l) for l in open(f)]
    tot={}
    for r in rows:
        print(f"{r['layer']:9s} {r['op']:6s} {r['cin']:5d} {r['cout']:5d} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
        t=tot.setdefault(r['op'],[0,0]); t[0]+=r['ours_ms']; t[1]+=r['cudnn_ms']
    print(tot)
PY
