#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_train.py -q -x > gpurun_out/tests27.log 2>&1; echo "exit $?" >> gpurun_out/tests27.log
timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --only L0.c1,L0.dec0,L1.c0,L1.c1,L1.dec0 --out gpurun_out/conv_layers27.jsonl > gpurun_out/conv_layers27.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench27.json 2> gpurun_out/bench27.err
B200SEG_ROWS2=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench27_norows2.json 2> gpurun_out/bench27_b.err
tail -5 gpurun_out/tests27.log; python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/conv_layers27.jsonl')]
for r in rows: print(f"{r['layer']:9s} {r['op']:6s} {r['ours_ms']:8.3f} {r['cudnn_ms']:8.3f} {r['ours_frac_of_measured_peak']:6.2f}")
PY
for f in bench27 bench27_norows2; do python -c "
import json
d=json.loads(open('gpurun_out/$f.json').read().strip().split('\n')[-1]);print('$f',d.get('value'),d.get('ms_per_step'))"; done; tail -3 gpurun_out/bench27.err
