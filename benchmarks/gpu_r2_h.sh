#!/bin/bash
mkdir -p gpurun_out
python benchmarks/profile_step.py --impl own --graph --steps 10 --out gpurun_out/r2h_step_graph.txt > gpurun_out/r2h_prof.log 2>&1; head -24 gpurun_out/r2h_step_graph.txt | cut -c1-140; tail -3 gpurun_out/r2h_prof.log | cut -c1-200
python benchmarks/profile_step.py --impl own --steps 10 --out gpurun_out/r2h_step_eager.txt > gpurun_out/r2h_prof2.log 2>&1; head -1 gpurun_out/r2h_step_eager.txt
