#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --workload turn_based --batch 8 --steps 5 --warmup 3 --repeats 2 > gpurun_out/r2_bench_turn_based_n2.json 2> gpurun_out/r2_bench_turn_based_n2.err
echo "tb n2 rc=$?"; tail -n 3 gpurun_out/r2_bench_turn_based_n2.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_turn_based_n2.json').read()); print(d['value'], d['ms_per_step'], d['graph']['enabled'], d.get('grad_sync'))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2_bench_n2_packed.json 2> gpurun_out/r2_bench_n2_packed.err
echo "pre n2 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_n2_packed.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d.get('grad_sync'))"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
