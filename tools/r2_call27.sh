#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -3
for m in 0 1 0 1; do echo "VT_GEMM_EPI16=$m"; VT_GEMM_PAIR=1 VT_GEMM_EPI16=$m timeout 300 python tools/kernel_cases.py --only "grad " 2>&1 | grep -E "^(gemm|dgrad)" ; VT_GEMM_PAIR=1 VT_GEMM_EPI16=$m timeout 300 python tools/kernel_cases.py --only "gemm " 2>&1 | grep -E "^(gemm|dgrad)"; done
