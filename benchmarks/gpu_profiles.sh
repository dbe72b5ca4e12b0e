#!/bin/bash
# refresh the per-kernel step profiles (both arms) and the per-layer conv table in gpurun_out/
mkdir -p gpurun_out
timeout 600 python benchmarks/profile_step.py --impl own --out gpurun_out/step_own.txt > gpurun_out/step_own.log 2>&1
timeout 600 python benchmarks/profile_step.py --impl reference --out gpurun_out/step_ref.txt > gpurun_out/step_ref.log 2>&1
timeout 900 python benchmarks/conv_layers.py --base 64 --out gpurun_out/conv_layers_b64.jsonl > gpurun_out/conv_layers_b64.log 2>&1
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/conv_layers_b64.jsonl')]
with open('gpurun_out/conv_layers_b64_n16_512.txt','w') as f:
    f.write("# per-layer conv timing, U-Net b=64, batch 16, 512x512 (CUDA events, 3 warm-up, 5 reps); ours = tcgen05 kernels (fprop includes the BN\n")
    f.write("# statistics epilogue, no conv bias, as in the engine), cudnn = the torch library call on the same tensors\n")
    f.write("# frac = ours TFLOP/s / measured sustained bf16 rate (MEASURED_PEAKS.json: 1405.9 TFLOP/s; the burst rate is 1656)\n")
    f.write("layer     op       HW   cin  cout  ours_ms  cudnn_ms  ours_TF   frac\n")
    tot={}
    for r in rows:
        f.write(f"{r['layer']:9s} {r['op']:6s} {r.get('HW',0):4d} {r.get('cin',0):5d} {r.get('cout',0):5d} {r['ours_ms']:8.3f} {r['cudnn_ms']:9.3f} {r.get('ours_tflops',0):8.1f} {r['ours_frac_of_measured_peak']:6.2f}\n")
        t=tot.setdefault(r['op'],[0,0]); t[0]+=r['ours_ms']; t[1]+=r['cudnn_ms']
    for k,(a,b) in tot.items(): f.write(f"# total {k}: ours {a:.2f} ms, cudnn {b:.2f} ms\n")
print(open('gpurun_out/conv_layers_b64_n16_512.txt').read()[-400:])
PY
head -5 gpurun_out/step_own.txt; head -3 gpurun_out/step_ref.txt
