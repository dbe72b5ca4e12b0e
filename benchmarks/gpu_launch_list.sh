mkdir -p gpurun_out
timeout 48 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_final.csv python benchmarks/ncu_target.py --steps 2 > gpurun_out/launches_final.log 2>&1; echo "ncu exit $?"; wc -l gpurun_out/launches_final.csv; tail -1 gpurun_out/launches_final.log
