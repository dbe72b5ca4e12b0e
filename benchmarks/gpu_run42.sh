#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests42.log 2>&1; echo "exit $?" >> gpurun_out/tests42.log
timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --only L0.stem,L0.c1,L0.dec0 --out gpurun_out/cl42.jsonl > /dev/null 2>&1
B200SEG_ROWS4=0 timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --only L0.stem,L0.c1 --out gpurun_out/cl42_off.jsonl > /dev/null 2>&1
for i in 1 2; do
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench42_on_$i.json 2> gpurun_out/bench42.err
B200SEG_ROWS4=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench42_off_$i.json 2> gpurun_out/bench42b.err
done
tail -8 gpurun_out/tests42.log
python - <<'PY'
import json
for f in ('cl42','cl42_off'):
    print(f)
    for l in open(f'gpurun_out/{f}.jsonl'):
        r=json.loads(l); print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
for f in ('on_1','off_1','on_2','off_2'):
    d=json.loads(open(f'gpurun_out/bench42_{f}.json').read().strip().split('\n')[-1]);print(f,d.get('value'),d.get('ms_per_step'))
PY
