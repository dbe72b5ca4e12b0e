#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x -k wgrad > gpurun_out/tests40.log 2>&1; echo "exit $?" >> gpurun_out/tests40.log
for base in 64 128; do
timeout 600 python benchmarks/conv_layers.py --base $base --ops wgrad --out gpurun_out/cl40_${base}_def.jsonl > /dev/null 2>&1
B200SEG_WGRAD_KERNEL=strip B200SEG_WGRAD_BN=128 timeout 600 python benchmarks/conv_layers.py --base $base --ops wgrad --out gpurun_out/cl40_${base}_s128.jsonl > /dev/null 2>&1
B200SEG_WGRAD_KERNEL=strip B200SEG_WGRAD_BN=64 timeout 600 python benchmarks/conv_layers.py --base $base --ops wgrad --out gpurun_out/cl40_${base}_s64.jsonl > /dev/null 2>&1
B200SEG_WGRAD_KERNEL=splitk timeout 600 python benchmarks/conv_layers.py --base $base --ops wgrad --out gpurun_out/cl40_${base}_sk.jsonl > /dev/null 2>&1
done
tail -3 gpurun_out/tests40.log
python - <<'PY'
import json
for base in (64,128):
    tabs={}
    for v in ('def','s128','s64','sk'):
        try:
            for l in open(f'gpurun_out/cl40_{base}_{v}.jsonl'):
                r=json.loads(l); tabs.setdefault(r['layer'],{})[v]=r['ours_ms']; tabs[r['layer']]['cudnn']=r['cudnn_ms']
        except Exception as e: print(v,e)
    print('base',base)
    for k,d in tabs.items(): print(f"{k:9s} "+' '.join(f"{v}={d.get(v,float('nan')):7.3f}" for v in ('def','s128','s64','sk','cudnn')))
PY
