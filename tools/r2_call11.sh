#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --repeats 1 --min-warm-s 0 --no-graph --no-roofline --no-cpu-baseline --no-gpu-eager --no-clocks"
timeout 300 $CMD > gpurun_out/plain_pre.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 1300 --csv --log-file gpurun_out/r2_launches_pretrain_b32_packed.csv $CMD > gpurun_out/ncu_pre.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/r2_launches_pretrain_b32_packed.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_fwd_tc|attn_bwd_fused|attn_dq_reduce|attn_bwd_prep" -s 40 -c 4 -o gpurun_out/r2_attn $CMD > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"; ls -la gpurun_out/r2_attn.ncu-rep
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_bench_b32_packed2.json 2>/dev/null; cut -c1-200 gpurun_out/r2_bench_b32_packed2.json
timeout 400 python bench.py --steps 20 --warmup 3 --dense --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_bench_b32_dense2.json 2>/dev/null; cut -c1-200 gpurun_out/r2_bench_b32_dense2.json
