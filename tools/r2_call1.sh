#!/bin/bash
# round 2, GPU call 1: the sync-free / graphed step -- tests, then bench lines (graph, eager, b=2)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu --maxfail=10 -q 2>&1 | tail -40 > gpurun_out/r2_tests1.log
echo "pytest rc=$?" >> gpurun_out/r2_tests1.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_b32.json 2> gpurun_out/r2_bench_b32.err
echo "bench rc=$?"
timeout 400 python bench.py --steps 20 --warmup 3 --batch 2 --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_bench_b2.json 2> gpurun_out/r2_bench_b2.err
timeout 400 python bench.py --steps 20 --warmup 3 --no-graph --no-cpu-baseline --no-gpu-eager > gpurun_out/r2_bench_b32_eager.json 2> gpurun_out/r2_bench_b32_eager.err
tail -5 gpurun_out/r2_tests1.log; for f in gpurun_out/r2_bench*.err; do echo "== $f"; tail -n 4 $f; done; cat gpurun_out/r2_bench_b32.json | cut -c1-1500
