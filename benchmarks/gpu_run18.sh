#!/bin/bash
mkdir -p gpurun_out
timeout 600 python benchmarks/profile_step.py --impl own --conv tcgen05 --out gpurun_out/prof18.txt > /dev/null 2> gpurun_out/prof18.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench18_a.json 2> gpurun_out/bench18_a.err
B200SEG_WGRAD_MT=2 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench18_mt2.json 2> gpurun_out/bench18_mt2.err
B200SEG_CONV_STRIP=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench18_nostrip.json 2> gpurun_out/bench18_nostrip.err
B200SEG_WGRAD_KERNEL=splitk timeout 600 python bench.py --steps 10 --warmup 3 --no-e2e > gpurun_out/bench18_splitk.json 2> gpurun_out/bench18_splitk.err
cut -c1-150 gpurun_out/prof18.txt | head -24
for f in a mt2 nostrip splitk; do python -c "
import json;d=json.loads(open('gpurun_out/bench18_$f.json').read().strip().split('\n')[-1]);print('$f',d['value'],d['ms_per_step'])"; done
