#!/bin/bash
for d in 0 1 2 4 8 5 12 13; do echo "VT_GEMM_DBG=$d"; VT_GEMM_PAIR=0 VT_GEMM_DBG=$d timeout 300 python tools/kernel_cases.py --only "gemm " 2>&1 | grep -E "^(gemm)" ; VT_GEMM_PAIR=0 VT_GEMM_DBG=$d timeout 300 python tools/kernel_cases.py --only "dgrad attn" 2>&1 | grep -E "^(dgrad)"; done
