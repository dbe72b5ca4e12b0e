#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests25.log 2>&1; echo "exit $?" >> gpurun_out/tests25.log
timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --out gpurun_out/conv_layers25.jsonl > gpurun_out/conv_layers25.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench25.json 2> gpurun_out/bench25.err
tail -5 gpurun_out/tests25.log; python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/conv_layers25.jsonl')]
for r in rows: print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
PY
python -c "
import json
d=json.loads(open('gpurun_out/bench25.json').read().strip().split('\n')[-1]);print(d.get('value'),d.get('ms_per_step'))"; tail -3 gpurun_out/bench25.err
