#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_packed_gpu.py tests/test_kernels_gpu.py -m gpu -q --maxfail=6 2>&1 | tail -60 > gpurun_out/r2_tests_packed.log
tail -25 gpurun_out/r2_tests_packed.log | cut -c1-300
