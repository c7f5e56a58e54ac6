#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
echo "pretrain n$N rc=$?"; tail -n 3 gpurun_out/r2_bench_n$N.err | cut -c1-300; cut -c1-420 gpurun_out/r2_bench_n$N.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload turn_based --steps 10 --warmup 3 > gpurun_out/r2_bench_turn_based_n$N.json 2> gpurun_out/r2_bench_turn_based_n$N.err
echo "turn_based n$N rc=$?"; tail -n 3 gpurun_out/r2_bench_turn_based_n$N.err | cut -c1-300; cut -c1-420 gpurun_out/r2_bench_turn_based_n$N.json
