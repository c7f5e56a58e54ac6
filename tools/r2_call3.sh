#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_agents_gpu.py -m gpu -q 2>&1 | tail -5 > gpurun_out/r2_tests_agents.log
for wk in viewpoint_select turn_based; do
  timeout 500 python bench.py --workload $wk --steps 10 --warmup 3 > gpurun_out/r2_bench_${wk}_b32.json 2> gpurun_out/r2_bench_${wk}_b32.err
  echo "$wk rc=$?"
done
timeout 300 python bench.py --workload viewpoint_select --batch 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_viewpoint_select_b8.json 2> gpurun_out/r2_bench_viewpoint_select_b8.err
timeout 300 python bench.py --workload turn_based --batch 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_turn_based_b4.json 2> gpurun_out/r2_bench_turn_based_b4.err
timeout 300 python bench.py --workload qa_classifier --steps 10 --warmup 3 > gpurun_out/r2_bench_qa_b32.json 2> gpurun_out/r2_bench_qa_b32.err
cat gpurun_out/r2_tests_agents.log
for f in gpurun_out/r2_bench_viewpoint*.err gpurun_out/r2_bench_turn*.err gpurun_out/r2_bench_qa*.err; do echo "== $f"; tail -n 5 $f | cut -c1-600; done
for f in gpurun_out/r2_bench_viewpoint*.json gpurun_out/r2_bench_turn*.json gpurun_out/r2_bench_qa*.json; do echo "== $f"; cut -c1-700 $f; done
