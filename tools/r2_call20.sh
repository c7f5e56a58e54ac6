#!/bin/bash
# same-box A/B of two library builds: tools/ab/lib_before.so vs the tree's (the python side is the tree's in both legs)
mkdir -p gpurun_out
L=visitron_b200/libvisitron_b200.so
cp $L /tmp/lib_after.so
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_packed_gpu.py -x -q -m gpu -k "attention or packed" 2>&1 | tail -3
timeout 300 python tools/kernel_cases.py --only "packed rows" 2>&1 | grep "attn " | tail -2
for which in after before after before; do
  if [ $which = before ]; then cp tools/ab/lib_before.so $L; else cp /tmp/lib_after.so $L; fi
  VT_NO_BWD_SCHED=$([ $which = before ] && echo 1 || echo 0) timeout 300 python bench.py --steps 20 --warmup 3 --no-gpu-eager > gpurun_out/ab_$which.json 2> gpurun_out/ab_$which.err
  python -c "
import json; d=json.loads(open('gpurun_out/ab_$which.json').read()); print('$which', d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['clocks']['sm_mhz'])" || tail -3 gpurun_out/ab_$which.err
done
cp /tmp/lib_after.so $L
