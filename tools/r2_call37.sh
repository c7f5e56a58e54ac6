#!/bin/bash
mkdir -p gpurun_out
timeout 1000 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for m in 0 -1 0 -1; do VT_GEMM_PAIR=$m timeout 300 python bench.py --steps 20 --warmup 3 --no-gpu-eager --no-cpu-baseline > gpurun_out/ab_pair_$m.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/ab_pair_$m.json').read()); print('pair=$m', d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"; done
