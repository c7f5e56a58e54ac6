#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_agents_gpu.py tests/test_agent_heads_gpu.py -m gpu --maxfail=12 -q 2>&1 | tail -150 > gpurun_out/r2_tests_agents.log
tail -30 gpurun_out/r2_tests_agents.log
