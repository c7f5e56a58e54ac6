#!/bin/bash
# parity report + kernel cases + ncu full set of the row kernels and the LSTM (one ncu run)
mkdir -p gpurun_out; rm -f gpurun_out/r2_parity.txt
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_graph_gpu.py -m gpu -q 2>&1 | tail -6
timeout 600 python tools/kernel_cases.py --json gpurun_out/r2_kernel_cases.json > gpurun_out/r2_kernel_cases.txt 2>&1; tail -14 gpurun_out/r2_kernel_cases.txt
CMD="python tools/kernel_cases.py --once"
timeout 300 $CMD > gpurun_out/plain_cases.log 2>&1 && timeout 1200 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:"ln_fwd_kernel|ln_bwd_kernel|colsum_partial|embed_ln|softmax_ce|adamw_kernel|lstm_seq" -o gpurun_out/r2_rows $CMD > gpurun_out/ncu_rows.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_rows.log; ls -la gpurun_out/r2_rows.ncu-rep
