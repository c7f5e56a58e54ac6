#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_packed_gpu.py -x -q -m gpu -k "attention or packed" 2>&1 | tail -3
PYTHONPATH=. timeout 120 python tools/keepbits_diag.py 2>&1 | grep "differing" | head -8
timeout 300 python tools/kernel_cases.py --only "attn bwd" 2>&1 | grep "attn " | tail -12
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_b32_v6_packed.json 2> gpurun_out/r2_bench_b32_v6.err
echo "bench rc=$?"; tail -n 3 gpurun_out/r2_bench_b32_v6.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_b32_v6_packed.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d['clocks'], d['roofline']['achieved'], d['roofline']['gemm_ms_per_step'], d['gpu_launches'])"
