#!/bin/bash
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -2
timeout 300 python tools/kernel_cases.py --only "grad " 2>&1 | grep -E "^(dgrad)"; timeout 300 python tools/kernel_cases.py --only "gemm " 2>&1 | grep -E "^(gemm)"
