#!/bin/bash
for d in 0 1 4 8 12 5; do echo "VT_GEMM_DBG=$d"; VT_GEMM_DBG=$d timeout 300 python tools/kernel_cases.py --only "wgrad" 2>&1 | grep -E "^(wgrad)"; done
