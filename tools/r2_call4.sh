#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload viewpoint_select --batch 8 --steps 2 --warmup 3 --repeats 1 --min-warm-s 0 --no-graph --no-roofline --no-cpu-baseline --no-clocks"
timeout 300 $CMD > gpurun_out/plain_vs.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 6000 --csv --log-file gpurun_out/r2_launches_vs_b8.csv $CMD > gpurun_out/ncu_vs.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_vs.log; wc -l gpurun_out/r2_launches_vs_b8.csv
