#!/bin/bash
mkdir -p gpurun_out
timeout 900 python benchmarks/conv_layers.py --base 128 --out gpurun_out/conv_layers39_b128.jsonl > gpurun_out/conv_layers39.log 2>&1
timeout 900 python benchmarks/conv_layers.py --base 64 --out gpurun_out/conv_layers39_b64.jsonl > gpurun_out/conv_layers39b.log 2>&1
timeout 600 python benchmarks/profile_step.py --impl own --base-channels 128 --out gpurun_out/step39_b128.txt > gpurun_out/step39.log 2>&1
python - <<'PY'
import json
for f in ('b128',):
    tot={}
    for l in open(f'gpurun_out/conv_layers39_{f}.jsonl'):
        r=json.loads(l); print(f"{r['layer']:9s} {r['op']:6s} {r['hw']:4d} {r['cin']:5d} {r['cout']:5d} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
        t=tot.setdefault(r['op'],[0,0]); t[0]+=r['ours_ms']; t[1]+=r['cudnn_ms']
    print(tot)
PY
head -16 gpurun_out/step39_b128.txt | cut -c1-110
