#!/bin/bash
mkdir -p gpurun_out
N=8
run() { name=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 10 --warmup 3 --repeats 3 --no-cpu-baseline --no-gpu-eager --no-roofline > gpurun_out/dp8_$name.json 2> gpurun_out/dp8_$name.err
  python -c "
import json; d=json.loads(open('gpurun_out/dp8_$name.json').read()); print('$name', d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'])" || tail -3 gpurun_out/dp8_$name.err
}
run skip_tail VT_DBG_SKIP_TAIL=1
run ctas8 VT_NCCL_MAX_CTAS=8
run ctas16 VT_NCCL_MAX_CTAS=16
grep -h -i "nvls" gpurun_out/dp8_*.err | head -3
