#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for lib in old new; do
  if [ $lib = old ]; then export B200SEG_LIB=$PWD/benchmarks/ab/lib_old.so; else unset B200SEG_LIB; fi
  timeout 600 python benchmarks/conv_layers.py --ops fprop,dgrad --only L2.dec0,L3.c1,L3.dec0,L4.dec0 --out gpurun_out/ab_${lib}_$rep.jsonl > /dev/null 2>&1
  timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/ab_bench_${lib}_$rep.json 2> /dev/null
done; done
python - <<'PY'
import json
for rep in (1,2):
  for lib in ('old','new'):
    rows=[json.loads(l) for l in open(f'gpurun_out/ab_{lib}_{rep}.jsonl')]
    d=json.loads(open(f'gpurun_out/ab_bench_{lib}_{rep}.json').read().strip().split('\n')[-1])
    print(lib, rep, ' '.join(f"{r['layer']}.{r['op'][0]}={r['ours_ms']:.3f}" for r in rows), 'bench', d['ms_per_step'])
PY
