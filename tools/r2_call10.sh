#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity.txt
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 2>&1 | tail -25 > gpurun_out/r2_tests_all.log
tail -12 gpurun_out/r2_tests_all.log | cut -c1-250
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_b32_packed.json 2> gpurun_out/r2_bench_b32_packed.err
echo "bench rc=$?"; tail -n 3 gpurun_out/r2_bench_b32_packed.err | cut -c1-300; cut -c1-900 gpurun_out/r2_bench_b32_packed.json
timeout 300 python bench.py --steps 20 --warmup 3 --batch 2 --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_bench_b2_packed.json 2>/dev/null; cut -c1-330 gpurun_out/r2_bench_b2_packed.json
