#!/bin/bash
mkdir -p gpurun_out
b() { tag=$1; shift; env "$@" timeout 600 python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/s_$tag.json 2> gpurun_out/s_$tag.err; python -c "
import json; d=json.loads(open('gpurun_out/s_$tag.json').read().strip().split('\n')[-1]); print('$tag', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"; }
b new X=1
b old B200SEG_STRIP_N=256 B200SEG_DGRAD_N64=0
b new2 X=1
b oldn B200SEG_STRIP_N=256
b oldd B200SEG_DGRAD_N64=0
