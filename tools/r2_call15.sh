#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -4
timeout 300 python tools/kernel_cases.py --only wgrad 2>&1 | grep -v Warn | tail -12
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_packed_gpu.py tests/test_graph_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_b32_v4_packed.json 2> gpurun_out/r2_bench_b32_v4.err
echo "bench rc=$?"; tail -n 3 gpurun_out/r2_bench_b32_v4.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_b32_v4_packed.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d['clocks'], d['roofline']['achieved'], d['roofline']['gemm_ms_per_step'], d['gpu_launches'])"
