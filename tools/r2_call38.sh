#!/bin/bash
# end-of-round profiles with the CTA-pair GEMMs: default bench line, per-kernel cases, ncu launch list
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r2_final3_b32.json 2> gpurun_out/r2_final3_b32.err; echo "default bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_final3_b32.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'], d['gpu_launches'], d['clocks'])"
timeout 400 python tools/kernel_cases.py > gpurun_out/r2_kernel_cases_v3.txt 2>&1; grep "layer total" gpurun_out/r2_kernel_cases_v3.txt
CMD="python bench.py --steps 2 --warmup 3 --repeats 1 --min-warm-s 0 --no-graph --no-roofline --no-cpu-baseline --no-gpu-eager --no-clocks"
timeout 300 $CMD > gpurun_out/plain_pre.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 1300 --csv --log-file gpurun_out/r2_launches_pretrain_b32_packed_v3.csv $CMD > gpurun_out/ncu_pre.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/r2_launches_pretrain_b32_packed_v3.csv
