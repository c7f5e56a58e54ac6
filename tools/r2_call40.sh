#!/bin/bash
for d in 0 1 2 4 12; do echo "VT_GEMM_DBG=$d"; VT_GEMM_DBG=$d timeout 300 python tools/kernel_cases.py --only "dgrad ffn-down" 2>&1 | grep -E "^(dgrad)"; VT_GEMM_DBG=$d timeout 300 python tools/kernel_cases.py --only "gemm qkv" 2>&1 | grep -E "^(gemm)"; done
