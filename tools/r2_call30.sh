#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline --no-gpu-eager --no-roofline > gpurun_out/dp_$name.json 2> gpurun_out/dp_$name.err
  python -c "
import json; d=json.loads(open('gpurun_out/dp_$name.json').read()); print('$name', d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'])" || tail -3 gpurun_out/dp_$name.err
}
run base X=1
run skip_tail VT_DBG_SKIP_TAIL=1
run no_reserve VT_DBG_NO_RESERVE=1
run skip_tail_no_reserve VT_DBG_SKIP_TAIL=1 VT_DBG_NO_RESERVE=1
run tail_capped VT_TAIL_FULL_SPEED=0
run ctas8 VT_NCCL_MAX_CTAS=8
