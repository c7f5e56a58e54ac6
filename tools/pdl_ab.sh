#!/bin/bash
# Same-box A/B of programmatic dependent launch (common.cuh launch_pdl): alternates VT_PDL=1 / VT_PDL=0 bench runs (result: profiles/r1_pdl_ab.txt).
#   usage: tools/pdl_ab.sh [rounds] [extra bench flags...]
R=${1:-3}; shift
for i in $(seq 1 $R); do
  for v in 1 0; do
    VT_PDL=$v python bench.py --no-cpu-baseline --no-roofline "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('VT_PDL=$v value %.3f ms %s  e2e %.2f ms  sm %s' % (d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['ms_per_step'], (d.get('clocks') or {}).get('sm_mhz')))"
  done
done
