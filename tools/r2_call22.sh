#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/kernel_cases.py --only "gemm ffn-up" 2>&1 | grep "gemm" | tail -3
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 --no-gpu-eager > gpurun_out/r2_bench_b32_v9_packed.json 2> gpurun_out/r2_bench_b32_v9.err
echo "bench rc=$?"; tail -n 3 gpurun_out/r2_bench_b32_v9.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_b32_v9_packed.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d['clocks'], d['roofline']['achieved'], d['roofline']['gemm_ms_per_step'], d['gpu_launches'])"
