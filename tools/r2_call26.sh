#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
# one full-set capture of the FFN-up GEMM (GELU + saved gelu' epilogue), single-CTA and pair
VT_GEMM_PAIR=0 timeout 200 python tools/kernel_cases.py --once --only "gemm ffn-up" > gpurun_out/kc_once.log 2>&1; echo "plain rc=$?"
VT_GEMM_PAIR=0 timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_bf16 -o gpurun_out/r2_gemm_ffn_up -f python tools/kernel_cases.py --once --only "gemm ffn-up" > gpurun_out/ncu_ffn.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_ffn.log
