#!/bin/bash
for w in 0 1 0 1; do echo "VT_LN_BWD_WIDE=$w"; VT_LN_BWD_WIDE=$w timeout 300 python tools/kernel_cases.py --only "ln bwd" 2>&1 | grep "ln bwd" | tail -2; done
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_packed_gpu.py tests/test_graph_gpu.py -x -q -m gpu 2>&1 | tail -3
for w in 0 1 0 1; do VT_LN_BWD_WIDE=$w timeout 300 python bench.py --steps 20 --warmup 3 --no-gpu-eager --no-cpu-baseline > gpurun_out/ab_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/ab_$w.json').read()); print('wide=$w', d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'])"; done
