#!/bin/bash
for f in 0 1024 2048 3072; do
echo "VT_ATTN_DBG=$f"; VT_ATTN_DBG=$f timeout 300 python tools/kernel_cases.py --only "reading the keep bits" 2>&1 | grep "attn " | tail -2
done
