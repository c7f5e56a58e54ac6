"""Kernel-level parity on the B200: every C-ABI entry point against a plain fp32 PyTorch statement of
the same op (the module-level parity against the oracle lives in test_parity_gpu.py).
Tolerances are bf16-level: inputs/outputs are bf16 with fp32 accumulation (bf16 eps = 2^-8)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from visitron_b200 import lib, ops

from oracle import visitron_oracle as vo

DEV = "cuda"
BF = torch.bfloat16


def close(a, b, rtol=2e-2, atol=2e-2, what=""):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    assert bool((err <= tol).all()), "%s: max err %.4g (ref absmax %.4g), %d / %d outside tolerance" % (
        what, float(err.max()), float(b.abs().max()), int((err > tol).sum()), err.numel())


def rel_l2(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


@pytest.mark.parametrize("a_major,b_major", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("bn", [0, 64, 128, 192, 256])
def test_gemm_majors(a_major, b_major, bn):
    torch.manual_seed(1)
    M, N, K = 304, 520, 264
    A = (torch.randn(M, K, device=DEV) * 0.5).to(BF)
    B = (torch.randn(N, K, device=DEV) * 0.5).to(BF)
    Ain = A.t().contiguous() if a_major else A
    Bin = B.t().contiguous() if b_major else B
    lib.init().vt_debug_force_block_n(bn)
    try:
        out = lib.gemm(Ain, Bin, a_major=a_major, b_major=b_major, out_dtype=torch.float32)
    finally:
        lib.init().vt_debug_force_block_n(0)
    close(out, A.float() @ B.float().t(), rtol=1e-3, atol=1e-3, what="gemm")


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 768, 264), (1000, 304, 3072), (5000, 2304, 768)])
@pytest.mark.parametrize("b_major", [0, 1])
def test_gemm_cta_pair_equals_single_cta(M, N, K, b_major):
    """tcgen05.mma.cta_group::2 (two SMs per 256 x 256 tile): same accumulation order per output element as the single-CTA kernel, so
    the results are bit-identical -- plain, with the fused epilogues, with a partial last row block and under a device row limit."""
    torch.manual_seed(M + K)
    A = (torch.randn(M, K, device=DEV) * 0.5).to(BF)
    B = (torch.randn(N, K, device=DEV) * 0.1).to(BF)
    Bin = B.t().contiguous() if b_major else B
    bias = torch.randn(N, device=DEV)
    res = torch.randn(M, N, device=DEV).to(BF)
    L = lib.init()
    outs = {}
    try:
        for mode in (0, 1):
            L.vt_debug_pair_mode(mode)
            o = [lib.gemm(A, Bin, b_major=b_major, out_dtype=torch.float32)]
            aux = torch.zeros(M, N, device=DEV, dtype=BF)
            o.append(lib.gemm(A, Bin, b_major=b_major, bias=bias, act=lib.ACT_GELU_SAVE_GRAD, aux_out=aux))
            o.append(aux)
            o.append(lib.gemm(A, Bin, b_major=b_major, bias=bias, residual=res, dropout_p=0.25, seed=5, dropout_stream=3))
            o.append(lib.gemm(A, Bin, b_major=b_major, act=lib.ACT_MUL_AUX, aux=res))
            if M >= 512:
                rows = M - 200
                lim = torch.tensor([(rows + 127) // 128 * 128, rows], device=DEV, dtype=torch.int32)
                buf = torch.full((M, N), 7.0, device=DEV, dtype=BF)
                with ops.row_limit(lim, 0):
                    lib.gemm(A, Bin, b_major=b_major, bias=bias, out=buf)
                assert bool((buf[(rows + 127) // 128 * 128:] == 7.0).all())      # rows past the limit are not written
                o.append(buf[:rows].clone())
            torch.cuda.synchronize()
            outs[mode] = o
    finally:
        L.vt_debug_pair_mode(-1)
    close(outs[1][0], A.float() @ B.float().t(), rtol=1e-3, atol=1e-3, what="pair gemm")
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("M,N,K", [(768, 768, 5000), (768, 3072, 9000), (520, 1000, 4100), (256, 256, 130)])
def test_gemm_weight_gradient_cta_pair(M, N, K):
    """dW = dY^T X with both operands MN-major on CTA pairs (256 x 256 tiles, split-K, fp32 red.add) against the single-CTA kernel and
    fp32 torch; also under a device row limit (the contraction stops at the row count)."""
    torch.manual_seed(M + K)
    dY = (torch.randn(K, M, device=DEV) * 0.2).to(BF)
    X = (torch.randn(K, N, device=DEV) * 0.2).to(BF)
    ref = dY.float().t() @ X.float()
    L = lib.init()
    outs = {}
    try:
        for mode in (0, 1):
            L.vt_debug_pair_mode(mode)
            acc = torch.full((M, N), 0.25, device=DEV)
            lib.gemm(dY, X, a_major=1, b_major=1, out=acc, accumulate=True, alpha=0.5)
            outs[mode] = acc
            if mode == 1 and K >= 512:
                rows = K - 300
                lim = torch.tensor([(rows + 127) // 128 * 128, rows], device=DEV, dtype=torch.int32)
                dZ = dY.clone()
                dZ[rows:] = 0
                acc2 = torch.zeros(M, N, device=DEV)
                with ops.row_limit(lim, 0):
                    lib.gemm(dZ, X, a_major=1, b_major=1, out=acc2, accumulate=True)
                close(acc2, dY[:rows].float().t() @ X[:rows].float(), rtol=1e-3, atol=3e-3, what="pair wgrad under a row limit")
    finally:
        L.vt_debug_pair_mode(-1)
    close(outs[1], 0.25 + 0.5 * ref, rtol=1e-3, atol=3e-3, what="pair wgrad")
    close(outs[1], outs[0], rtol=1e-4, atol=1e-3, what="pair vs single wgrad")


def test_gemm_epilogues_and_splitk():
    torch.manual_seed(2)
    M, N, K = 640, 768, 768
    A = (torch.randn(M, K, device=DEV) * 0.5).to(BF)
    B = (torch.randn(N, K, device=DEV) * 0.05).to(BF)
    bias = torch.randn(N, device=DEV)
    res = torch.randn(M, N, device=DEV).to(BF)
    base = A.float() @ B.float().t() + bias
    aux = torch.empty(M, N, device=DEV, dtype=BF)
    close(lib.gemm(A, B, bias=bias, act=lib.ACT_GELU, aux_out=aux), torch.nn.functional.gelu(base), what="gelu")
    close(aux, base, what="aux_out")
    close(lib.gemm(A, B, bias=bias, residual=res), base + res.float(), what="residual")
    close(lib.gemm(A, B, bias=bias, act=lib.ACT_TANH, out_dtype=torch.float32), torch.tanh(base), rtol=1e-3, atol=1e-3, what="tanh")
    pre = torch.randn(M, N, device=DEV).to(BF)
    x = pre.float().requires_grad_(True)
    torch.nn.functional.gelu(x).sum().backward()
    close(lib.gemm(A, B, act=lib.ACT_GELU_BWD, aux=pre), (A.float() @ B.float().t()) * x.grad, what="gelu_bwd")
    # the FFN pair: forward saves gelu'(x) from the fp32 accumulator, backward multiplies by it
    dg = torch.empty_like(aux)
    close(lib.gemm(A, B, bias=bias, act=lib.ACT_GELU_SAVE_GRAD, aux_out=dg), torch.nn.functional.gelu(base), what="gelu (save grad)")
    xb = base.detach().clone().requires_grad_(True)
    torch.nn.functional.gelu(xb).sum().backward()
    close(dg, xb.grad, what="saved gelu'")
    close(lib.gemm(A, B, act=lib.ACT_MUL_AUX, aux=dg), (A.float() @ B.float().t()) * dg.float(), what="mul aux")
    # weight-gradient shape: both operands MN-major, split-K with fp32 atomics, accumulating into a live buffer
    dY = (torch.randn(4096, 768, device=DEV) * 0.1).to(BF)
    X = (torch.randn(4096, 3072, device=DEV) * 0.1).to(BF)
    acc = torch.ones(768, 3072, device=DEV)
    lib.gemm(dY, X, a_major=1, b_major=1, out=acc, accumulate=True)
    close(acc, 1.0 + dY.float().t() @ X.float(), rtol=1e-3, atol=1e-3, what="wgrad split-K")
    # alpha from the device
    g = torch.tensor([0.25], device=DEV)
    close(lib.gemm(A, B, alpha=2.0, alpha_dev=g, out_dtype=torch.float32), 0.5 * (A.float() @ B.float().t()), rtol=1e-3, atol=1e-3, what="alpha")
    # odd N (vocab-like) with fp32 output and a padded leading dimension
    Bv = (torch.randn(3001, K, device=DEV) * 0.05).to(BF)
    out = torch.zeros(M, 3008, device=DEV)
    lib.gemm(A, Bv, out=out[:, :3001], bias=torch.zeros(3001, device=DEV))
    close(out[:, :3001], A.float() @ Bv.float().t(), rtol=1e-3, atol=1e-3, what="odd N")
    assert float(out[:, 3001:].abs().max()) == 0.0


@pytest.mark.parametrize("M,N,K,bn", [(2304, 768, 5000, 0), (3072, 768, 17500, 0), (3001, 768, 130, 0), (100, 200, 64, 0),
                                      (768, 768, 32, 0), (520, 304, 1000, 64), (520, 304, 1000, 128), (520, 384, 1000, 192)])
def test_gemm_weight_gradient_with_bias_gradient(M, N, K, bn):
    """colsum: the bias gradient (sum over rows of dY) as a side product of dW = dY^T X; also under split-K, on partial row blocks,
    with a padded leading dimension and under a device row limit (only the first `rows` rows of dY / X count)."""
    torch.manual_seed(M + K)
    ld = (M + 7) // 8 * 8 + 8
    dY = torch.zeros(K, ld, device=DEV, dtype=BF)
    dY[:, :M] = (torch.randn(K, M, device=DEV) * 0.2 + 0.05).to(BF)
    X = (torch.randn(K, N, device=DEV) * 0.2).to(BF)
    acc = torch.full((M, N), 0.5, device=DEV)
    db = torch.full((M + 3,), 2.0, device=DEV)
    lib.init().vt_debug_force_block_n(bn)
    try:
        lib.gemm(dY[:, :M], X, a_major=1, b_major=1, out=acc, accumulate=True, colsum=db)
        close(acc, 0.5 + dY[:, :M].float().t() @ X.float(), rtol=1e-3, atol=2e-3, what="dW")
        close(db[:M], 2.0 + dY[:, :M].float().sum(0), rtol=1e-3, atol=2e-3, what="db")
        assert bool((db[M:] == 2.0).all())
        if K >= 256:
            rows = K - 200
            lim = torch.tensor([(rows + 127) // 128 * 128, rows], device=DEV, dtype=torch.int32)
            dY[rows:] = 0          # the pad-row rule: rows past the count are finite zeros (the last 64-row block is contracted whole)
            acc.fill_(0.0)
            db.fill_(0.0)
            with ops.row_limit(lim, 0):
                lib.gemm(dY[:, :M], X, a_major=1, b_major=1, out=acc, accumulate=True, colsum=db, alpha=0.5)
            close(acc, 0.5 * (dY[:rows, :M].float().t() @ X[:rows].float()), rtol=1e-3, atol=2e-3, what="dW limited")
            close(db[:M], 0.5 * dY[:rows, :M].float().sum(0), rtol=1e-3, atol=2e-3, what="db limited")
    finally:
        lib.init().vt_debug_force_block_n(0)
    with pytest.raises(lib.VtError):
        lib.gemm(X, X, out_dtype=torch.float32, colsum=db)      # K-major A: not the weight-gradient form


def test_gemm_row_remap_and_strided_a():
    torch.manual_seed(3)
    b, R, S, T = 4, 24, 64, 40
    A = (torch.randn(b * R, 136, device=DEV)).to(BF)
    W = (torch.randn(768, 136, device=DEV) * 0.1).to(BF)
    out = torch.zeros(b * S, 768, device=DEV, dtype=BF)
    lib.gemm(A, W, out=out, remap=(R, S, T))
    ref = (A.float() @ W.float().t()).view(b, R, 768)
    close(out.view(b, S, 768)[:, T:], ref, what="remap")
    assert float(out.view(b, S, 768)[:, :T].abs().max()) == 0.0
    # A = row 0 of every sample (pooler): a strided view, no copy
    seq = torch.randn(b, S, 768, device=DEV).to(BF)
    first = seq.view(b * S, 768)[::S]
    Wp = (torch.randn(768, 768, device=DEV) * 0.05).to(BF)
    close(lib.gemm(first, Wp, out_dtype=torch.float32), seq[:, 0].float() @ Wp.float().t(), rtol=1e-3, atol=1e-3, what="strided A")


@pytest.mark.parametrize("rows", [1, 37, 5000])
def test_layernorm_fwd_bwd(rows):
    torch.manual_seed(4)
    x = (torch.randn(rows, 768, device=DEV) * 2 + 0.3).to(BF)
    gamma = 1 + 0.1 * torch.randn(768, device=DEV)
    beta = 0.1 * torch.randn(768, device=DEV)
    dy = torch.randn(rows, 768, device=DEV).to(BF)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    yr = vo.bert_layer_norm(xr, gr, br, 1e-12)
    yr.backward(dy.float())
    y = ops.layernorm_fwd(x, gamma, beta, 1e-12)
    close(y, yr, what="ln fwd")
    dgamma, dbeta, dbias = torch.zeros(768, device=DEV), torch.zeros(768, device=DEV), torch.zeros(768, device=DEV)
    dx, dxd = ops.layernorm_bwd(dy, x, gamma, 1e-12, dgamma, dbeta, dbias)
    assert dxd is dx
    close(dx, xr.grad, rtol=2e-2, atol=2e-2 * float(xr.grad.abs().max()), what="ln dx")
    assert rel_l2(dgamma, gr.grad) < 5e-3 and rel_l2(dbeta, br.grad) < 5e-3
    assert rel_l2(dbias, dx.float().sum(0)) < 5e-3


def test_dropout_mask_consistency_gemm_vs_ln_bwd():
    """The mask the GEMM epilogue applied in forward must be the one LayerNorm-backward regenerates."""
    torch.manual_seed(5)
    M = 1000
    A = torch.randn(M, 768, device=DEV).to(BF)
    W = (torch.randn(768, 768, device=DEV) * 0.05).to(BF)
    p, seed, sid = 0.3, 1234, 17
    h_drop = lib.gemm(A, W, dropout_p=p, seed=seed, dropout_stream=sid)
    h = lib.gemm(A, W)
    kept = h_drop != 0
    nz = h != 0
    frac = float(kept[nz].float().mean())
    assert abs(frac - 0.7) < 0.005, frac
    close(h_drop[kept], h[kept].float() / (1 - round(p * 65536) / 65536), what="kept values scaled")
    x = torch.randn(M, 768, device=DEV).to(BF)
    dy = (torch.randn(M, 768, device=DEV) + 3).to(BF)
    gamma = torch.ones(768, device=DEV)
    z = torch.zeros(768, device=DEV)
    dx, dxd = ops.layernorm_bwd(dy, x, gamma, 1e-12, z.clone(), z.clone(), z.clone(), out_dropout_p=p, out_stream=sid, seed=seed)
    nzd = dx != 0
    assert bool(((dxd != 0)[nzd & nz] == kept[nzd & nz]).all())
    # and a different stream id gives a different mask
    _, dxd2 = ops.layernorm_bwd(dy, x, gamma, 1e-12, z.clone(), z.clone(), z.clone(), out_dropout_p=p, out_stream=sid + 1, seed=seed)
    assert float(((dxd2 != 0) == (dxd != 0)).float().mean()) < 0.7


@pytest.mark.parametrize("with_types,with_pos", [(True, False), (False, False), (True, True)])
def test_embeddings_fwd_bwd(with_types, with_pos):
    torch.manual_seed(6)
    b, T, S, V, P = 3, 50, 70, 1000, 64
    word = (0.05 * torch.randn(V, 768, device=DEV)).requires_grad_(True)
    pos = (0.05 * torch.randn(P, 768, device=DEV)).requires_grad_(True)
    typ = (0.05 * torch.randn(2, 768, device=DEV)).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(768, device=DEV)).requires_grad_(True)
    beta = (0.1 * torch.randn(768, device=DEV)).requires_grad_(True)
    ids = torch.randint(0, V, (b, T), device=DEV)
    ids[:, 5:9] = 103  # repeated ids -> colliding scatter-adds
    tt = torch.randint(0, 2, (b, T), device=DEV) if with_types else None
    pp = torch.randint(0, P, (b, T), device=DEV) if with_pos else None
    out = torch.zeros(b * S, 768, device=DEV, dtype=BF)
    ops.embed_ln_fwd(ids, tt, pp, word.detach(), pos.detach(), typ.detach(), gamma.detach(), beta.detach(), 1e-12, out, S)
    P_ = {"bert.embeddings.word_embeddings.weight": word, "bert.embeddings.position_embeddings.weight": pos,
          "bert.embeddings.token_type_embeddings.weight": typ, "bert.embeddings.LayerNorm.weight": gamma,
          "bert.embeddings.LayerNorm.bias": beta}
    ref = vo.embeddings(P_, vo.OracleConfig(), ids, tt, pp)
    close(out.view(b, S, 768)[:, :T], ref, what="embed fwd")
    dy = torch.randn(b, S, 768, device=DEV).to(BF)
    ref.backward(dy[:, :T].float())
    g = [torch.zeros_like(t) for t in (word, pos, typ, gamma, beta)]
    ops.embed_ln_bwd(dy.view(b * S, 768), ids, tt, pp, word.detach(), pos.detach(), typ.detach(), gamma.detach(), 1e-12, S, *g)
    for name, mine, ref_t in zip(("word", "pos", "type", "gamma", "beta"), g, (word, pos, typ, gamma, beta)):
        assert rel_l2(mine, ref_t.grad) < 1e-2, (name, rel_l2(mine, ref_t.grad))


def test_small_row_ops():
    torch.manual_seed(7)
    x = torch.randn(3000, 2304, device=DEV).to(BF)
    out = torch.ones(2304, device=DEV)
    ops.colsum(x, out)
    assert rel_l2(out - 1, x.float().sum(0)) < 1e-3
    xv = torch.randn(77, 30528, device=DEV).to(BF)
    outv = torch.zeros(30522, device=DEV)
    ops.colsum(xv, outv, ncols=30522, accumulate=False)
    assert rel_l2(outv, xv[:, :30522].float().sum(0)) < 1e-3
    w = torch.randn(700, 2054, device=DEV)
    d = torch.zeros(700, 2184, device=DEV, dtype=BF)
    dt = torch.zeros(2054, 704, device=DEV, dtype=BF)
    ops.cast_weight(w, d[:, :2054], dt[:, :700])
    assert torch.equal(d[:, :2054], w.to(BF)) and torch.equal(dt[:, :700], w.t().to(BF))
    assert float(d[:, 2054:].abs().max()) == 0
    a, b2 = torch.rand(50, 2054, device=DEV), torch.rand(50, 128, device=DEV)
    o = torch.full((50, 2184), 7.0, device=DEV, dtype=BF)
    ops.pad_cast_concat(a, b2, o)
    assert torch.equal(o[:, :2054], a.to(BF)) and torch.equal(o[:, 2054:2182], b2.to(BF)) and float(o[:, 2182:].abs().max()) == 0
    src = torch.randn(500, 768, device=DEV).to(BF)
    idx = torch.randperm(500, device=DEV)[:123]
    gth = ops.gather_rows(src, idx)
    assert torch.equal(gth, src[idx])
    dst = torch.randn(500, 768, device=DEV).to(BF)
    ref = dst.clone()
    ref[idx] = (ref[idx].float() + gth.float()).to(BF)
    ops.scatter_add_rows(gth, idx, dst)
    assert torch.equal(dst, ref)
    m = torch.tensor([[1, 0, 255, 254]], dtype=torch.uint8, device=DEV)
    assert ops.mask_bias(m).tolist() == [[0.0, -10000.0, 2540000.0, 2530000.0]]
    assert ops.mask_bias(m.long()).tolist() == [[0.0, -10000.0, 2540000.0, 2530000.0]]
    f = torch.randn(64, 768, device=DEV)
    assert torch.equal(ops.to_f32(f.to(BF)), f.to(BF).float())
    assert torch.equal(ops.to_bf16(f), f.to(BF))


def _attn_ref(qkv, bias, B, S, heads, mask=None, scale_drop=1.0):
    q, k, v = [t.view(B, S, heads, 64).permute(0, 2, 1, 3) for t in qkv.view(B * S, 3, heads * 64).unbind(1)]
    s = q @ k.transpose(-1, -2) / 8.0 + bias[:, None, None, :]
    p = torch.softmax(s, -1)
    if mask is not None:
        p = p * mask * scale_drop
    return (p @ v).permute(0, 2, 1, 3).reshape(B * S, heads * 64)


@pytest.mark.parametrize("B,S,heads,mask_kind", [(2, 767, 12, "pad"), (3, 511, 12, "f6"), (1, 64, 2, "pad"), (2, 100, 3, "none"), (1, 1, 1, "none"),
                                                 (1, 129, 1, "pad"), (2, 1000, 2, "pad"),
                                                 (9, 767, 12, "pad")])   # enough items for two query tiles per CTA in the forward
def test_attention_fwd_bwd(B, S, heads, mask_kind):
    torch.manual_seed(8)
    Hh = heads * 64
    qkv = torch.randn(B * S, 3 * Hh, device=DEV).to(BF)
    if mask_kind == "pad":
        mask = torch.ones(B, S, dtype=torch.long, device=DEV)
        mask[0, S // 2:] = 0
        mask[-1, S - 3:] = 0
    elif mask_kind == "f6":
        mask = ~torch.zeros(B, S, dtype=torch.uint8, device=DEV)
    else:
        mask = torch.ones(B, S, dtype=torch.long, device=DEV)
    bias = ops.mask_bias(mask)
    close(bias, vo.extended_mask(mask, torch.float32)[:, 0, 0], rtol=0, atol=0, what="bias")
    ctx, lse = ops.attention_fwd(qkv, bias, B, S, heads)
    qr = qkv.float().requires_grad_(True)
    ref = _attn_ref(qr, bias, B, S, heads)
    close(ctx, ref, what="attention fwd")
    dctx = torch.randn(B * S, Hh, device=DEV).to(BF)
    ref.backward(dctx.float())
    dqkv = ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads)
    scale = float(qr.grad.abs().max())
    close(dqkv, qr.grad, rtol=3e-2, atol=2e-2 * scale, what="attention bwd")
    assert rel_l2(dqkv, qr.grad) < 2e-2


def test_attention_dropout_fwd_bwd_same_mask():
    """Recover the dropout mask with one-hot V, then check fwd and bwd against autograd with that mask."""
    torch.manual_seed(9)
    B, S, heads, p, seed, sid = 2, 128, 2, 0.1, 99, 5
    Hh = heads * 64
    qkv = torch.randn(B * S, 3 * Hh, device=DEV).to(BF)
    bias = torch.zeros(B, S, device=DEV)
    mask = torch.zeros(B, heads, S, S, device=DEV)
    for blk in range(S // 64):
        probe = qkv.clone().view(B, S, 3, heads, 64)
        probe[:, :, 2] = 0
        for j in range(64):
            probe[:, blk * 64 + j, 2, :, j] = 1
        o, _ = ops.attention_fwd(probe.view(B * S, 3 * Hh), bias, B, S, heads, dropout_p=p, seed=seed, stream_id=sid)
        mask[:, :, :, blk * 64:(blk + 1) * 64] = (o.view(B, S, heads, 64).permute(0, 2, 1, 3) != 0).float()
    keep = float(mask.mean())
    assert abs(keep - 0.9) < 0.01, keep
    sd = 1.0 / (1.0 - round(p * 65536) / 65536)
    ctx, lse = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=seed, stream_id=sid)
    qr = qkv.float().requires_grad_(True)
    ref = _attn_ref(qr, bias, B, S, heads, mask, sd)
    close(ctx, ref, what="dropout attention fwd")
    dctx = torch.randn(B * S, Hh, device=DEV).to(BF)
    ref.backward(dctx.float())
    dqkv = ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=p, seed=seed, stream_id=sid)
    assert rel_l2(dqkv, qr.grad) < 2e-2, rel_l2(dqkv, qr.grad)
    # the stored keep bits are the generator's decisions: the bits agree with the recovered mask, the forward's output does not
    # change, and the backward that reads them returns the very same bits as the one that draws them again
    bits = ops.attention_keep_bits(B, S, heads, DEV)
    bits.fill_(0x55555555)
    ctx_b, lse_b = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=seed, stream_id=sid, keep_bits=bits)
    assert torch.equal(ctx_b, ctx) and torch.equal(lse_b, lse)
    G = 4 * ((S + 127) // 128)
    w = bits.view(B * heads, G, S).long() & 0xffffffff
    unpacked = ((w[:, :, :, None] >> torch.arange(32, device=DEV)) & 1).permute(0, 2, 1, 3).reshape(B, heads, S, G * 32)[..., :S]
    assert torch.equal(unpacked.float(), mask)
    dqkv_b = ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=p, seed=seed, stream_id=sid, keep_bits=bits)
    assert int((dqkv_b != dqkv).sum()) <= 1e-4 * dqkv.numel() and rel_l2(dqkv_b, dqkv) < 1e-4     # a stray ulp at most (see test_packed_gpu.py)
    bits.zero_()      # ... and it really reads them: with every probability dropped only the -P.rowsum(dO.O) term is left
    dqkv_0 = ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=p, seed=seed, stream_id=sid, keep_bits=bits)
    assert not torch.equal(dqkv_0, dqkv) and float(dqkv_0[:, 2 * Hh:].abs().max()) == 0.0


@pytest.mark.parametrize("ncols,double", [(30522, False), (1601, True), (96, True), (5, False)])
def test_softmax_ce(ncols, double):
    torch.manual_seed(10)
    rows, ld = 33, ops.roundup(ncols, 8) + 8
    logits = torch.randn(rows, ld, device=DEV) * 3
    labels = torch.randint(0, ncols, (rows,), device=DEV)
    row_loss, argmax, stats = ops.softmax_ce_fwd(logits[:, :ncols], ncols, labels, double)
    x = logits[:, :ncols].clone().requires_grad_(True)
    inp = torch.softmax(x, -1) if double else x
    ref = torch.nn.functional.cross_entropy(inp, labels, reduction="none")
    close(row_loss, ref, rtol=1e-4, atol=1e-5, what="ce loss")
    assert torch.equal(argmax, x.argmax(1))
    g = torch.tensor([0.7], device=DEV)
    (ref.sum() * 0.7 / rows).backward()
    d = ops.softmax_ce_bwd(logits[:, :ncols], ncols, labels, stats, 1.0 / rows, g, double)
    assert rel_l2(d[:, :ncols], x.grad) < 6e-3
    assert float(d[:, ncols:].abs().max()) == 0 if d.size(1) > ncols else True


def test_action_head_and_finalize():
    torch.manual_seed(11)
    B, A = 13, 36
    pooled = torch.tanh(torch.randn(B, 768, device=DEV))
    W = (0.05 * torch.randn(A, 768, device=DEV))
    bias = 0.05 * torch.randn(A, device=DEV)
    labels = torch.randint(0, A, (B,), device=DEV)
    labels[3] = -1
    pr, Wr, br = pooled.clone().requires_grad_(True), W.clone().requires_grad_(True), bias.clone().requires_grad_(True)
    sc = torch.log_softmax(torch.nn.functional.linear(pr, Wr, br), -1)
    loss = torch.nn.functional.cross_entropy(sc, labels, ignore_index=-1)
    (loss * 0.5).backward()
    scores, row_loss, argmax = ops.action_head_fwd(pooled, W, bias, labels)
    close(scores, sc, rtol=1e-4, atol=1e-5, what="action scores")
    assert torch.equal(argmax, sc.argmax(1))
    fake = (torch.zeros(0, device=DEV), torch.zeros(0, dtype=torch.long, device=DEV), torch.zeros(0, dtype=torch.long, device=DEV))
    out = ops.pretrain_finalize(fake, fake, (row_loss, argmax, labels))
    assert abs(float(out[2]) - float(loss)) < 1e-5 and float(out[7]) == B - 1
    assert abs(float(out[5]) - float((sc.argmax(1) == labels).float().mean())) < 1e-6
    dW, db = torch.zeros_like(W), torch.zeros_like(bias)
    dp = ops.action_head_bwd(pooled, W, labels, scores, out[7:8], torch.tensor([0.5], device=DEV), dW, db)
    assert rel_l2(dp, pr.grad) < 1e-4 and rel_l2(dW, Wr.grad) < 1e-4 and rel_l2(db, br.grad) < 1e-4
    # mlm / token slots of finalize
    rl = torch.rand(7, device=DEV)
    am = torch.tensor([1, 2, 3, 4, 5, 6, 7], device=DEV)
    lb = torch.tensor([1, 2, 0, 4, 0, 6, 0], device=DEV)
    out = ops.pretrain_finalize((rl, am, lb), (rl * 2, am, am), (row_loss, argmax, labels))
    assert abs(float(out[1]) - float(rl.mean())) < 1e-6 and abs(float(out[3]) - 2 * float(rl.mean())) < 1e-6
    assert abs(float(out[4]) - 4 / 7) < 1e-6 and float(out[6]) == 1.0
    assert abs(float(out[0]) - float(out[1] + out[2] + out[3])) < 1e-6
    y = torch.tanh(torch.randn(4, 768, device=DEV))
    dy = torch.randn(4, 768, device=DEV)
    close(ops.tanh_bwd(dy, y), dy * (1 - y * y), what="tanh bwd")


@pytest.mark.parametrize("B,S,heads,p", [(3, 767, 2, 0.0), (4, 767, 3, 0.1), (2, 511, 2, 0.1), (2, 100, 1, 0.0), (2, 2048, 1, 0.1), (9, 767, 12, 0.1)])
def test_attention_key_tile_plan_changes_no_bit(B, S, heads, p):
    """Skipping the key tiles that are masked as a whole (vt_attention_key_tiles) must leave ctx, the saved row statistics and every
    gradient bit-identical to the kernels that visit all tiles: masked keys have probability exactly 0 (exp(-10000 + ...) underflows)."""
    g = torch.Generator().manual_seed(B * S)
    mask = torch.zeros(B, S, dtype=torch.long)
    T = min(511, S)
    for b in range(B):                       # [text | pad | regions | pad] like the loader's batches; sample 0 has a one-token text
        lt = 1 if b == 0 else int(torch.randint(T // 5, T + 1, (1,), generator=g))
        mask[b, :lt] = 1
        if S > T:
            mask[b, T:T + int(torch.randint(1, S - T + 1, (1,), generator=g))] = 1
    if B > 2:
        mask[2] = 0                          # a sample with EVERY key masked: uniform softmax over all keys, every tile stays live
    mask = mask.to(DEV)
    qkv = (torch.randn(B * S, 3 * heads * 64, generator=g) * 0.7).to(DEV).to(torch.bfloat16)
    dctx = (torch.randn(B * S, heads * 64, generator=g)).to(DEV).to(torch.bfloat16)
    bias = ops.mask_bias(mask)
    plan = ops.attention_key_tiles(bias)
    nkt = (S + 63) // 64
    want = []
    for b in range(B):
        m = mask[b].cpu()
        bits = sum(1 << t for t in range(nkt) if bool(m[t * 64:(t + 1) * 64].any()))
        want.append(bits if bits else (1 << nkt) - 1)
    assert [int(x) & 0xffffffff for x in plan.tolist()] == want
    assert any(bin(w).count("1") < nkt for w in want), "the test must skip something"
    ctx0, lse0 = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=7, stream_id=3)
    ctx1, lse1 = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=7, stream_id=3, key_tiles=plan)
    assert torch.equal(ctx0, ctx1) and torch.equal(lse0, lse1)
    d0 = ops.attention_bwd(qkv, bias, ctx0, dctx, lse0, B, S, heads, dropout_p=p, seed=7, stream_id=3)
    d1 = ops.attention_bwd(qkv, bias, ctx0, dctx, lse0, B, S, heads, dropout_p=p, seed=7, stream_id=3, key_tiles=plan)
    assert torch.equal(d0, d1)


def test_attention_key_tile_plan_with_fractional_masks():
    """Mask values other than 0 / 1 (bias = (1 - m) * -10000, vs/encoder.py:226-241): a tile is dead only when ALL its biases are at
    least 5000 below the sample's largest; anything closer is visited.  Same bits with and without the plan."""
    B, S, heads = 2, 511, 2
    g = torch.Generator().manual_seed(5)
    mask = torch.ones(B, S)
    mask[0, 64:128] = 0.5      # bias -5000: dead (probabilities underflow to exactly 0)
    mask[0, 128:192] = 0.6     # bias -4000: live by the rule
    mask[0, 192:256] = 0.0
    mask[1, :] = 0.25          # every key at -7500: the largest bias of the sample, all tiles live
    mask[1, 300:] = -0.5       # bias -15000: 7500 below -> dead
    mask = mask.to(DEV)
    qkv = (torch.randn(B * S, 3 * heads * 64, generator=g) * 0.7).to(DEV).to(torch.bfloat16)
    bias = ops.mask_bias(mask)
    plan = [int(x) & 0xffffffff for x in ops.attention_key_tiles(bias).tolist()]
    assert plan[0] == 0b11110101 and plan[1] == 0b00011111, [bin(x) for x in plan]
    pl = ops.attention_key_tiles(bias)
    for p in (0.0, 0.1):
        a = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=3, stream_id=1)
        b = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=p, seed=3, stream_id=1, key_tiles=pl)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_attention_with_plan_stays_inside_ctx_and_dqkv():
    """Guard bands around ctx and dqkv (no compute-sanitizer on the pool): sentinel rows before and behind stay untouched."""
    B, S, heads = 3, 767, 2
    Hh = heads * 64
    g = torch.Generator().manual_seed(1)
    mask = torch.ones(B, S, dtype=torch.long)
    mask[0, 70:511] = 0
    mask[1, 600:] = 0
    mask = mask.to(DEV)
    qkv = (torch.randn(B * S, 3 * Hh, generator=g) * 0.7).to(DEV).to(torch.bfloat16)
    dctx = torch.randn(B * S, Hh, generator=g).to(DEV).to(torch.bfloat16)
    bias = ops.mask_bias(mask)
    plan = ops.attention_key_tiles(bias)
    G = 64
    big = torch.full((B * S + 2 * G, Hh), 7.0, dtype=torch.bfloat16, device=DEV)
    ctx, lse = ops.attention_fwd(qkv, bias, B, S, heads, dropout_p=0.1, seed=2, stream_id=0, ctx=big[G:G + B * S], key_tiles=plan)
    bigd = torch.full((B * S + 2 * G, 3 * Hh), 7.0, dtype=torch.bfloat16, device=DEV)
    ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=0.1, seed=2, stream_id=0, dqkv=bigd[G:G + B * S], key_tiles=plan)
    torch.cuda.synchronize()
    for t in (big, bigd):
        assert bool((t[:G] == 7.0).all()) and bool((t[G + B * S:] == 7.0).all())
    assert bool(torch.isfinite(ctx.float()).all()) and bool(torch.isfinite(bigd[G:G + B * S].float()).all())
