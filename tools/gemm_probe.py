"""Diagnostic sweep of the tcgen05 GEMM against torch.matmul (run on the GPU box)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from visitron_b200 import lib

torch.manual_seed(0)
dev = "cuda"
L = lib.init()
res = []

def run(name, M, N, K, a_major, b_major, bn=0, **kw):
    L.vt_debug_force_block_n(bn)
    A = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.5).bfloat16()
    Ain = A.t().contiguous() if a_major else A
    Bin = B.t().contiguous() if b_major else B
    ref = A.float() @ B.float().t()
    out_dtype = kw.pop("out_dtype", torch.bfloat16)
    try:
        if kw.get("accumulate"):
            out = torch.zeros(M, N, device=dev, dtype=torch.float32)
            lib.gemm(Ain, Bin, a_major=a_major, b_major=b_major, out=out, **kw)
        else:
            out = lib.gemm(Ain, Bin, a_major=a_major, b_major=b_major, out_dtype=out_dtype, **kw)
        torch.cuda.synchronize()
    except Exception as e:
        res.append(dict(name=name, error=str(e)))
        print(name, "ERROR", e, flush=True)
        return
    err = (out.float() - ref).abs()
    tol = 0.02 * ref.abs() + 0.05
    bad = (err > tol)
    r = dict(name=name, M=M, N=N, K=K, a=a_major, b=b_major, bn=bn, max_err=float(err.max()), ref_absmax=float(ref.abs().max()),
             bad_frac=float(bad.float().mean()))
    if bad.any():
        rows = bad.any(1).nonzero().flatten()[:8].tolist()
        cols = bad.any(0).nonzero().flatten()[:8].tolist()
        r.update(bad_rows=rows, bad_cols=cols, sample_out=out.float()[:2, :4].tolist(), sample_ref=ref[:2, :4].tolist())
    res.append(r)
    print(json.dumps(r), flush=True)

# the four operand-major combinations, every tile width, ragged edges
for bn in (0, 64, 128, 192, 256):
    run("nt", 256, 512, 256, 0, 0, bn)
run("nt_k64", 128, 256, 64, 0, 0, 256)
run("nt_ragged", 300, 200, 136, 0, 0)
run("nt_big", 4096, 768, 768, 0, 0)
run("nt_ffn", 4096, 3072, 768, 0, 0)
run("nt_ffn2", 4096, 768, 3072, 0, 0)
for bn in (64, 128, 192, 256):
    run("dgrad(K,MN)", 384, 768, 512, 0, 1, bn)
    run("wgrad(MN,MN)", 768, 768, 1024, 1, 1, bn, out_dtype=torch.float32)
run("A_MN only", 256, 256, 256, 1, 0, 128)
run("wgrad_splitk", 768, 768, 8192, 1, 1, 0, accumulate=True)
run("wgrad_splitk_ffn", 3072, 768, 4096, 1, 1, 0, accumulate=True)
run("vocab_like", 200, 30522, 768, 0, 0, out_dtype=torch.float32)
run("f32out", 256, 256, 128, 0, 0, out_dtype=torch.float32)

# epilogues
L.vt_debug_force_block_n(0)
M, N, K = 512, 768, 768
A = (torch.randn(M, K, device=dev) * 0.5).bfloat16(); B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
bias = torch.randn(N, device=dev); resid = torch.randn(M, N, device=dev).bfloat16()
base = A.float() @ B.float().t() + bias
aux = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
o = lib.gemm(A, B, bias=bias, act=lib.ACT_GELU, aux_out=aux)
print("gelu", float((o.float() - torch.nn.functional.gelu(base)).abs().max()), "aux", float((aux.float() - base).abs().max()))
o = lib.gemm(A, B, bias=bias, residual=resid)
print("bias+resid", float((o.float() - (base + resid.float())).abs().max()))
o = lib.gemm(A, B, bias=bias, act=lib.ACT_TANH)
print("tanh", float((o.float() - torch.tanh(base)).abs().max()))
pre = torch.randn(M, N, device=dev).bfloat16()
o = lib.gemm(A, B, act=lib.ACT_GELU_BWD, aux=pre)
x = pre.float().requires_grad_(True); torch.nn.functional.gelu(x).sum().backward()
print("gelu_bwd", float((o.float() - (A.float() @ B.float().t()) * x.grad).abs().max()))
o1 = lib.gemm(A, B, bias=bias, dropout_p=0.3, seed=5, dropout_stream=1)
o2 = lib.gemm(A, B, bias=bias, dropout_p=0.3, seed=5, dropout_stream=1)
keep = (o1 != 0).float().mean()
print("dropout keep", float(keep), "deterministic", bool(torch.equal(o1, o2)), "kept vals err", float(((o1.float() - base / 0.7) * (o1 != 0)).abs().max()))
big = torch.zeros(4 * 96, N, device=dev, dtype=torch.bfloat16)
lib.gemm(A[:4 * 32], B, out=big, bias=bias, remap=(32, 96, 64))
chk = big.view(4, 96, N)[:, 64:, :].reshape(-1, N)
print("remap", float((chk.float() - base[:128]).abs().max()), "untouched", float(big.view(4, 96, N)[:, :64].abs().max()))

# timing of the encoder shapes
def bench(M, N, K, a=0, b=0, iters=20, **kw):
    A = torch.randn((K, M) if a else (M, K), device=dev).bfloat16(); B = torch.randn((K, N) if b else (N, K), device=dev).bfloat16()
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if kw.get("accumulate") else torch.bfloat16)
    for _ in range(3): lib.gemm(A, B, a_major=a, b_major=b, out=out, **kw)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): lib.gemm(A, B, a_major=a, b_major=b, out=out, **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    tf = 2.0 * M * N * K / ms / 1e9
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    Bm = B.t().contiguous() if b else B; Am = A.t().contiguous() if a else A
    for _ in range(3): torch.matmul(Am, Bm.t())
    t0.record()
    for _ in range(iters): torch.matmul(Am, Bm.t())
    t1.record(); torch.cuda.synchronize()
    print(f"time M={M} N={N} K={K} a={a} b={b} {kw}: {ms:.3f} ms {tf:.0f} TFLOP/s | cublas {2.0*M*N*K/(t0.elapsed_time(t1)/iters)/1e9:.0f}", flush=True)

Mrows = 32 * 767
bench(Mrows, 2304, 768); bench(Mrows, 768, 768); bench(Mrows, 3072, 768); bench(Mrows, 768, 3072)
bench(Mrows, 768, 2304, 0, 1); bench(Mrows, 3072, 768, 0, 1)
bench(768, 768, Mrows, 1, 1, accumulate=True); bench(3072, 768, Mrows, 1, 1, accumulate=True); bench(2304, 768, Mrows, 1, 1, accumulate=True)
json.dump(res, open("gpurun_out/gemm_probe.json", "w"), indent=1)
