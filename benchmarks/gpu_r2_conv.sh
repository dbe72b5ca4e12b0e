#!/bin/bash
# conv kernels: correctness tests, then the per-layer table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py -q -x > gpurun_out/r2_conv_tests.log 2>&1; echo "exit $?" >> gpurun_out/r2_conv_tests.log
tail -3 gpurun_out/r2_conv_tests.log
python benchmarks/conv_layers.py --reps 5 --out gpurun_out/r2_conv_layers_${1:-a}.jsonl > gpurun_out/r2_conv_layers_${1:-a}.log 2>&1
python - <<PY
import json
for l in open('gpurun_out/r2_conv_layers_${1:-a}.jsonl'):
    r=json.loads(l); print(f"{r['layer']:8s} {r['op']:6s} ours {r['ours_ms']:.3f} cudnn {r['cudnn_ms']:.3f} frac {r['ours_frac_of_measured_peak']:.2f}")
PY
tail -1 gpurun_out/r2_conv_layers_${1:-a}.log
