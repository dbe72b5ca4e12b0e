#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests17.log 2>&1; echo "exit $?" >> gpurun_out/tests17.log
timeout 900 python benchmarks/conv_layers.py --ops fprop,dgrad --out gpurun_out/conv_layers17.jsonl > gpurun_out/conv_layers17.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench17.json 2> gpurun_out/bench17.err
tail -3 gpurun_out/tests17.log; python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/conv_layers17.jsonl')]
for r in rows: print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
PY
cat gpurun_out/bench17.json; tail -3 gpurun_out/bench17.err
