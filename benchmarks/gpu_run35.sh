#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests35.log 2>&1; echo "exit $?" >> gpurun_out/tests35.log
timeout 600 python benchmarks/conv_layers.py --ops wgrad --out gpurun_out/conv_layers35.jsonl > gpurun_out/conv_layers35.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench35.json 2> gpurun_out/bench35.err
B200SEG_CONVT_QUADS=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench35_noquads.json 2> gpurun_out/bench35_b.err
timeout 600 python benchmarks/profile_step.py --impl own --out gpurun_out/step35.txt > gpurun_out/step35.log 2>&1
tail -4 gpurun_out/tests35.log; python - <<'PY'
import json
for l in open('gpurun_out/conv_layers35.jsonl'):
    r=json.loads(l); print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
for f in ('bench35','bench35_noquads'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'))
PY
grep -E "wgrad|total" gpurun_out/step35.txt | cut -c1-110
