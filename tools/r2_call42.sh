#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/kernel_cases.py --once --only "dgrad ffn-down" > gpurun_out/kc_once.log 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_bf16 -o gpurun_out/r2_gemm_dgrad_ffn_down -f python tools/kernel_cases.py --once --only "dgrad ffn-down" > gpurun_out/ncu_dg.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_dg.log
