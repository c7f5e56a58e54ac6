for d in 0 511; do echo -n "dbg=$d "; VT_ATTN_DBG=$d python tools/kernel_cases.py --only "attn" 2>&1 | grep "attn [fb]" | cut -c1-8,60- | tr "\n" " "; echo; done
