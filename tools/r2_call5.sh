#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_agents_gpu.py tests/test_agent_heads_gpu.py -m gpu -q 2>&1 | tail -30 > gpurun_out/r2_tests_agents.log
tail -5 gpurun_out/r2_tests_agents.log
for cfg in "viewpoint_select 8" "viewpoint_select 32" "turn_based 4" "turn_based 32"; do
  set -- $cfg
  timeout 400 python bench.py --workload $1 --batch $2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_$1_b$2.json 2> gpurun_out/r2_bench_$1_b$2.err
  echo "$1 b$2 rc=$?"; tail -n 2 gpurun_out/r2_bench_$1_b$2.err | cut -c1-300; cut -c1-330 gpurun_out/r2_bench_$1_b$2.json
done
