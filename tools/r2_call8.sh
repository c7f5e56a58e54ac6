#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_agents_gpu.py -m gpu -q 2>&1 | tail -3
timeout 200 python tools/kernel_cases.py --only lstm 2>&1 | tail -2
for b in 1 2 4 8 16 32 64 128 256 512; do
  timeout 200 python bench.py --workload qa_classifier --batch $b --steps 5 --warmup 3 --repeats 3 --no-roofline --no-cpu-baseline --no-clocks 2> gpurun_out/qa_$b.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('qa_classifier b=%d  %.1f samples/s  %.3f ms/step  e2e %.1f  graph=%s eager_ms=%s' % (d['config']['per_gpu_batch'], d['value'], d['ms_per_step'], d['e2e']['value'], d['graph']['enabled'], d['graph']['eager_ms_per_step']))" >> gpurun_out/r2_qa_sweep.txt
done
cat gpurun_out/r2_qa_sweep.txt
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 --cpu-batch 8 > gpurun_out/r2_reference_arm_b8.json 2>/dev/null; cut -c1-400 gpurun_out/r2_reference_arm_b8.json
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_reference_arm_b2.json 2>/dev/null; cut -c1-300 gpurun_out/r2_reference_arm_b2.json
