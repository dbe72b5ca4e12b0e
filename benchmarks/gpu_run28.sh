#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests28.log 2>&1; echo "exit $?" >> gpurun_out/tests28.log
timeout 600 python benchmarks/conv_layers.py --ops wgrad --out gpurun_out/conv_layers28.jsonl > gpurun_out/conv_layers28.log 2>&1
B200SEG_WGRAD_BN=64 timeout 600 python benchmarks/conv_layers.py --ops wgrad --only L1.c0,L1.c1,L1.dec0,L2.c0,L2.c1,L2.dec0,L3.c1,L3.dec0 --out gpurun_out/conv_layers28_bn64.jsonl > gpurun_out/conv_layers28b.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench28.json 2> gpurun_out/bench28.err
B200SEG_WGRAD_BN=64 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench28_bn64.json 2> gpurun_out/bench28_b.err
tail -5 gpurun_out/tests28.log; python - <<'PY'
import json
for f in ('conv_layers28','conv_layers28_bn64'):
    print(f)
    for l in open(f'gpurun_out/{f}.jsonl'):
        r=json.loads(l); print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
for f in ('bench28','bench28_bn64'):
    d=json.loads(open(f'gpurun_out/{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'))
PY
tail -3 gpurun_out/bench28.err
