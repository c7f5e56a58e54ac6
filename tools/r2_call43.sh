#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -2
timeout 300 python tools/kernel_cases.py --only "grad " 2>&1 | grep -E "^(dgrad|wgrad)"; timeout 300 python tools/kernel_cases.py --only "gemm " 2>&1 | grep -E "^(gemm)"
timeout 300 python bench.py --steps 20 --warmup 3 --no-gpu-eager --no-cpu-baseline > gpurun_out/ab_x.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/ab_x.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'], d['roofline']['achieved'])"
