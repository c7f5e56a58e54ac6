#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
echo "n2 rc=$?"; tail -n 6 gpurun_out/r2_bench_n2.err | cut -c1-500; cut -c1-2500 gpurun_out/r2_bench_n2.json
