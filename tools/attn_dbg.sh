for d in 0 1 2 4 7 8 16 32 40 47 63; do echo -n "dbg=$d "; VT_ATTN_DBG=$d python tools/kernel_cases.py --only "attn bwd" 2>&1 | grep "attn bwd" | awk '{print $(NF-3), $(NF-2)}'; done
