#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/kernel_cases.py --once --only "reading the keep bits" > gpurun_out/kc_once.log 2>&1
echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attn_bwd_fused -o gpurun_out/r2_attn_bwd_bits -f python tools/kernel_cases.py --once --only "reading the keep bits" > gpurun_out/ncu_bits.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_bits.log
ls -la gpurun_out/*.ncu-rep
