#!/bin/bash
L=visitron_b200/libvisitron_b200.so
cp $L /tmp/lib_after.so
for which in before after before after; do
  if [ $which = before ]; then cp tools/ab/lib_before.so $L; else cp /tmp/lib_after.so $L; fi
  echo "== $which"; timeout 300 python tools/kernel_cases.py --only "attn fwd" 2>&1 | grep "attn fwd" | tail -4
done
cp /tmp/lib_after.so $L
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_packed_gpu.py -x -q -m gpu -k "attention or packed" 2>&1 | tail -3
