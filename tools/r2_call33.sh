#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2_bench_n${N}_v3.json 2> gpurun_out/r2_bench_n${N}_v3.err
echo "pretrain n$N rc=$?"; tail -n 2 gpurun_out/r2_bench_n${N}_v3.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_n${N}_v3.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d.get('grad_sync'), d['clocks'])"
