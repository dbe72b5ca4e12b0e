#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests23.log 2>&1; echo "exit $?" >> gpurun_out/tests23.log
timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --only L0.c1,L0.dec0,L1.c0,L1.c1,L1.dec0,L2.c1 --out gpurun_out/conv_layers23.jsonl > gpurun_out/conv_layers23.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench23.json 2> gpurun_out/bench23.err
B200SEG_CLUSTER=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench23_nocl.json 2> gpurun_out/bench23_nocl.err
tail -5 gpurun_out/tests23.log; python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/conv_layers23.jsonl')]
for r in rows: print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
PY
for f in gpurun_out/bench23.json gpurun_out/bench23_nocl.json; do python -c "
import json
d=json.loads(open('$f').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'))"; done; tail -3 gpurun_out/bench23.err
