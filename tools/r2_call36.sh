#!/bin/bash
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -3
for m in 0 -1 0 -1; do echo "VT_GEMM_PAIR=$m"; VT_GEMM_PAIR=$m timeout 300 python tools/kernel_cases.py --only "wgrad" 2>&1 | grep -E "^(wgrad)"; done
