#!/bin/bash
# end-of-round validation on one GPU: build check, smoke, the whole GPU suite, the default bench line, the reference arm
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -2
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r2_final2_b32.json 2> gpurun_out/r2_final2_b32.err; echo "default bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_final2_b32.json').read()); print(d['value'], d['ms_per_step'], d['timed_regions']['ms_per_step'], d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'], d['gpu_launches'], d['clocks'], d['cpu_baseline']['value'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final2_reference_arm.json 2>/dev/null; echo "reference arm rc=$?"; cut -c1-300 gpurun_out/r2_final2_reference_arm.json
