#!/bin/bash
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_packed_gpu.py -x -q -m gpu -k "attention or packed" 2>&1 | tail -3
timeout 300 python tools/kernel_cases.py --only "attn bwd" 2>&1 | grep "attn bwd"
timeout 300 python tools/kernel_cases.py --only "attn bwd" 2>&1 | grep "attn bwd"
