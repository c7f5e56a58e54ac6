# needs an ablation build:  VT_BUILD_ABLATE=1 python -m visitron_b200.build --force   (and a plain --force build afterwards)
# timing ablations of the tcgen05 attention kernels (results of dbg != 0 are garbage).  VT_ATTN_DBG bits: 1 no element-wise arithmetic,
# 2 no exponentials, 16 no dV (fwd: no P.V) MMA, 32 no dK MMA, 64 no dQ MMA, 128 no S / dP MMA, 256 no dQ-partial stores,
# 512 no K/V (Q/dO) loads after the first ring fill, 1024 no maximum exchange, 2048 no proxy fence, 4096 no P stores, 8192 no TMEM loads of S,
# 16384 two threads per query row in the forward, 32768 persistent forward grid
for d in ${DBGS:-0 1 17 49 113 241 497 129 256}; do echo -n "dbg=$d "; VT_ATTN_DBG=$d python tools/kernel_cases.py --only "attn" --p-attn ${PA:-0.1} 2>&1 | grep "attn [fb]" | cut -c1-8,60- | tr "\n" " "; echo; done
