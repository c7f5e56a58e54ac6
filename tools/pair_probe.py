"""CTA-pair GEMM (tcgen05.mma.cta_group::2): correctness against fp32 torch and timing against the single-CTA kernel."""
import sys
import torch
from visitron_b200 import lib
DEV, BF = "cuda", torch.bfloat16
L = lib.init()
def check(M, N, K, b_major, **kw):
    torch.manual_seed(M + N + K)
    A = (torch.randn(M, K, device=DEV) * 0.5).to(BF)
    B = (torch.randn(N, K, device=DEV) * 0.5).to(BF)
    Bin = B.t().contiguous() if b_major else B
    ref = A.float() @ B.float().t()
    out = {}
    for mode in (0, 1):
        L.vt_debug_pair_mode(mode)
        out[mode] = lib.gemm(A, Bin, b_major=b_major, out_dtype=torch.float32, **kw)
        torch.cuda.synchronize()
    L.vt_debug_pair_mode(-1)
    e0 = float((out[0] - ref).abs().max()); e1 = float((out[1] - ref).abs().max())
    print("M %5d N %5d K %5d b_major %d: max err single %.3g pair %.3g  equal %s" % (M, N, K, b_major, e0, e1, bool(torch.equal(out[0], out[1]))), flush=True)
    assert e1 <= 1e-3 + 1e-3 * float(ref.abs().max()), "pair GEMM wrong"
for M, N, K in [(256, 256, 64), (256, 256, 256), (512, 768, 768), (304, 520, 264), (1000, 304, 3072), (24544, 768, 768)]:
    for bm in (0, 1):
        check(M, N, K, bm)
if "--time" in sys.argv:
    def timeit(fn, n=20):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / n
    M = 24544
    for N, K, bm in [(2304, 768, 0), (768, 768, 0), (3072, 768, 0), (768, 3072, 0), (3072, 768, 1), (768, 3072, 1), (768, 2304, 1), (768, 768, 1)]:
        As = [(torch.randn(M, K, device=DEV) * 0.5).to(BF) for _ in range(4)]
        B = (torch.randn(N, K, device=DEV) * 0.05).to(BF)
        Bin = B.t().contiguous() if bm else B
        out = torch.empty(M, N, device=DEV, dtype=BF)
        r = []
        for mode in (0, 1, 0, 1):
            L.vt_debug_pair_mode(mode)
            i = [0]
            def fn():
                i[0] += 1
                lib.gemm(As[i[0] % 4], Bin, b_major=bm, out=out)
            r.append(timeit(fn))
        L.vt_debug_pair_mode(-1)
        fl = 2.0 * M * N * K
        print("N %4d K %4d b_major %d: single %.1f / %.1f us (%.0f TF/s)   pair %.1f / %.1f us (%.0f TF/s)" % (N, K, bm, r[0], r[2], fl / min(r[0], r[2]) / 1e6, r[1], r[3], fl / min(r[1], r[3]) / 1e6), flush=True)
