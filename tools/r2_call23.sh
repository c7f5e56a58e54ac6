#!/bin/bash
# round-2 final single-GPU measurements
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/r2_final_b32.json 2> gpurun_out/r2_final_b32.err; echo "default rc=$?"; cut -c1-300 gpurun_out/r2_final_b32.json
timeout 300 python bench.py --dense --steps 10 --warmup 3 --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_final_b32_dense.json 2>/dev/null; cut -c1-200 gpurun_out/r2_final_b32_dense.json
timeout 300 python bench.py --batch 2 --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_final_b2.json 2>/dev/null; cut -c1-200 gpurun_out/r2_final_b2.json
for wl in viewpoint_select turn_based qa_classifier; do
  timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_final_$wl.json 2>/dev/null; cut -c1-200 gpurun_out/r2_final_$wl.json
done
timeout 300 python bench.py --workload viewpoint_select --batch 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_final_viewpoint_select_b8.json 2>/dev/null; cut -c1-200 gpurun_out/r2_final_viewpoint_select_b8.json
timeout 300 python bench.py --workload turn_based --batch 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_final_turn_based_b4.json 2>/dev/null; cut -c1-200 gpurun_out/r2_final_turn_based_b4.json
timeout 400 python tools/kernel_cases.py > gpurun_out/r2_kernel_cases_v2.txt 2>&1; tail -5 gpurun_out/r2_kernel_cases_v2.txt
CMD="python bench.py --steps 2 --warmup 3 --repeats 1 --min-warm-s 0 --no-graph --no-roofline --no-cpu-baseline --no-gpu-eager --no-clocks"
timeout 300 $CMD > gpurun_out/plain_pre.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 1300 --csv --log-file gpurun_out/r2_launches_pretrain_b32_packed_v2.csv $CMD > gpurun_out/ncu_pre.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/r2_launches_pretrain_b32_packed_v2.csv
