# BASELINE configs[4] (large-batch sweep, GEMM vs attention vs launch-bound crossover) and the per-GPU batch sweep of the headline workload.
#   bash tools/batch_sweep.sh > profiles/rN_batch_sweep.txt      (one B200; ~2 minutes)
for w in qa_classifier pretrain viewpoint_select; do
  for b in 2 8 16 32 64; do
    python bench.py --workload $w --batch $b --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
r=d['roofline']; m=d['model_flops']
print('%-16s b=%-3d %9.1f samples/s %8.3f ms/step  e2e %9.1f  GEMM %6.1f TFLOP/s (%.0f%% of step)  executed model FLOPs %5.1f%% of peak  launches/step %d' % (
  '$w', $b, d['value'], d['ms_per_step'], d['e2e']['value'], r['achieved'], 100*r['gemm_share_of_step'], 100*m['frac_of_peak_executed'], d['gpu_launches']/d['steps']))"
  done
done
