#!/bin/bash
# Diagnosis: is the device-timed region (`value`) disturbed by the in-process NVML clock sampler?  Alternates runs with and
# without it on the SAME box and prints value / e2e ms per step.   usage: tools/bench_noise.sh [rounds] > gpurun_out/noise.txt
R=${1:-3}
for i in $(seq 1 $R); do
  for flag in "" "--no-clocks"; do
    python bench.py --no-cpu-baseline --no-roofline $flag 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('clocks=%-5s value %.2f ms  e2e %.2f ms  sm %s samples %s' % ('$flag'=='' , d['ms_per_step'], d['e2e']['ms_per_step'], (d.get('clocks') or {}).get('sm_mhz'), (d.get('clocks') or {}).get('samples')))"
  done
done
