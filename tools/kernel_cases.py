"""Per-kernel timing of ONE encoder layer (forward + backward) at the bench shape, through the C ABI.

    python tools/kernel_cases.py [--batch 32] [--seq 767] [--iters 10] [--once] [--json out.json]

Each case is timed with CUDA events on the launching stream (3 warm-ups, `iters` back-to-back launches, inputs far larger
than L2 are rotated so nothing is served from cache) and reported against its own roofline: TFLOP/s for the tcgen05
kernels, GB/s of algorithmic bytes for the row kernels.  `--once` launches every case once untimed, then once between cudaProfilerStart/Stop (for `ncu --set full
--profile-from-start off`).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from visitron_b200 import lib, ops
from visitron_b200.lib import ACT_GELU_SAVE_GRAD, ACT_MUL_AUX

H, I, NH = 768, 3072, 12
BF16, F32 = torch.bfloat16, torch.float32


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=767)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--p-hidden", type=float, default=0.3)
    ap.add_argument("--p-attn", type=float, default=0.1)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib.init()
    torch.manual_seed(0)
    B, S = args.batch, args.seq
    M = B * S
    ph, pa, seed = args.p_hidden, args.p_attn, 1234
    NROT = 3  # rotate activations so that consecutive launches do not hit L2

    def rnd(*shape, scale=1.0, dtype=BF16):
        return (torch.randn(*shape, device=dev) * scale).to(dtype)

    xs = [rnd(M, H) for _ in range(NROT)]
    zs = [rnd(M, I) for _ in range(NROT)]
    qkvs = [rnd(M, 3 * H) for _ in range(NROT)]
    Wqkv, Wo, W1, W2 = rnd(3 * H, H, scale=0.02), rnd(H, H, scale=0.02), rnd(I, H, scale=0.02), rnd(H, I, scale=0.02)
    bqkv, bo, b1, b2 = (torch.zeros(n, device=dev) for n in (3 * H, H, I, H))
    gamma, beta = torch.ones(H, device=dev), torch.zeros(H, device=dev)
    mask = torch.ones(B, S, dtype=torch.int64, device=dev)
    mask[:, S - 76:] = 0
    bias = ops.mask_bias(mask)
    out_h = torch.empty(M, H, dtype=BF16, device=dev)
    out_h2 = torch.empty(M, H, dtype=BF16, device=dev)
    out_i = torch.empty(M, I, dtype=BF16, device=dev)
    out_i2 = torch.empty(M, I, dtype=BF16, device=dev)
    out_q = torch.empty(M, 3 * H, dtype=BF16, device=dev)
    gW = torch.zeros(I * H, dtype=F32, device=dev)
    gvec = torch.zeros(4 * I, dtype=F32, device=dev)
    ctx0, lse0 = ops.attention_fwd(qkvs[0], bias, B, S, NH, pa, seed, 0)
    # the bench's own padding ([text 40-100 % of 511 | 180 of 256 regions]) and its key-tile plan
    from visitron_b200 import synthetic
    bmask = synthetic.pretrain_batch(B, 88)["attention_mask"].to(dev) if S == 767 else mask
    bbias = ops.mask_bias(bmask)
    plan = ops.attention_key_tiles(bbias)
    ctx1, lse1 = ops.attention_fwd(qkvs[0], bbias, B, S, NH, pa, seed, 0, key_tiles=plan)
    torch.cuda.synchronize()

    gemm_fl = lambda n, k: 2.0 * M * n * k
    attn_f = 4.0 * B * NH * S * S * 64
    cases = []

    def case(name, fn, flops=0.0, bytes_=0.0):
        cases.append((name, fn, flops, bytes_))

    r = lambda L, i: L[i % NROT]
    case("gemm qkv         [M,768]x[2304,768]  bias", lambda i: lib.gemm(r(xs, i), Wqkv, bias=bqkv, out=out_q), gemm_fl(3 * H, H))
    case("attn fwd         S=%d d=64 p=%.1f" % (S, pa), lambda i: ops.attention_fwd(r(qkvs, i), bias, B, S, NH, pa, seed, 0, ctx=out_h), attn_f)
    case("attn fwd         bench padding, masked key tiles skipped", lambda i: ops.attention_fwd(r(qkvs, i), bbias, B, S, NH, pa, seed, 0, ctx=out_h, key_tiles=plan), attn_f)
    kbits = ops.attention_keep_bits(B, S, NH, dev) if pa > 0 else None
    if kbits is not None:
        ops.attention_fwd(qkvs[0], bbias, B, S, NH, pa, seed, 0, key_tiles=plan, keep_bits=kbits)
        case("attn fwd         ... storing the keep bits", lambda i: ops.attention_fwd(r(qkvs, i), bbias, B, S, NH, pa, seed, 0, ctx=out_h, key_tiles=plan, keep_bits=kbits), attn_f)
    case("gemm attn-out    [M,768]x[768,768]   bias+dropout+residual",
         lambda i: lib.gemm(r(xs, i), Wo, bias=bo, dropout_p=ph, seed=seed, dropout_stream=1, residual=r(xs, i + 1), out=out_h), gemm_fl(H, H))
    case("ln fwd           [M,768]", lambda i: ops.layernorm_fwd(r(xs, i), gamma, beta, 1e-12, out=out_h), 0, 2.0 * M * H * 2)
    case("gemm ffn-up      [M,768]x[3072,768]  bias+gelu, saves gelu'", lambda i: lib.gemm(r(xs, i), W1, bias=b1, act=ACT_GELU_SAVE_GRAD, aux_out=out_i2, out=out_i), gemm_fl(I, H))
    case("gemm ffn-down    [M,3072]x[768,3072] bias+dropout+residual",
         lambda i: lib.gemm(r(zs, i), W2, bias=b2, dropout_p=ph, seed=seed, dropout_stream=2, residual=r(xs, i), out=out_h), gemm_fl(H, I))
    case("ln bwd           [M,768] (+dropped copy, dgamma/dbeta/dbias)",
         lambda i: ops.layernorm_bwd(r(xs, i), r(xs, i + 1), gamma, 1e-12, gvec[:H], gvec[H:2 * H], gvec[2 * H:3 * H], out_dropout_p=ph, out_stream=2,
                                     seed=seed, dx=out_h, dxd=out_h2), 0, 4.0 * M * H * 2)
    case("ln bwd           [M,768] (+dropped copy, dgamma/dbeta; bias gradient left to the wgrad GEMM)",
         lambda i: ops.layernorm_bwd(r(xs, i), r(xs, i + 1), gamma, 1e-12, gvec[:H], gvec[H:2 * H], None, out_dropout_p=ph, out_stream=2,
                                     seed=seed, dx=out_h, dxd=out_h2), 0, 4.0 * M * H * 2)
    case("wgrad ffn-down   dW[768,3072] = dY^T Z (CTA pairs, bias gradient from the LayerNorm backward)", lambda i: lib.gemm(r(xs, i), r(zs, i), a_major=1, b_major=1, out=gW.view(H, I), accumulate=True), gemm_fl(I, H))
    case("wgrad ffn-down   dW[768,3072] = dY^T Z (+ bias gradient)", lambda i: lib.gemm(r(xs, i), r(zs, i), a_major=1, b_major=1, out=gW.view(H, I), accumulate=True, colsum=gvec[:H]), gemm_fl(I, H))
    case("dgrad ffn-down   [M,768]x[768,3072]  x saved gelu'", lambda i: lib.gemm(r(xs, i), W2, b_major=1, act=ACT_MUL_AUX, aux=r(zs, i), out=out_i), gemm_fl(I, H))
    case("colsum           [M,3072] (separate kernel, no longer in the step)", lambda i: ops.colsum(r(zs, i), gvec[:I]), 0, 1.0 * M * I * 2)
    case("wgrad ffn-up     dW[3072,768] = dPre^T Y", lambda i: lib.gemm(r(zs, i), r(xs, i), a_major=1, b_major=1, out=gW.view(I, H), accumulate=True), gemm_fl(I, H))
    case("wgrad ffn-up     + bias gradient (column sums of dPre)", lambda i: lib.gemm(r(zs, i), r(xs, i), a_major=1, b_major=1, out=gW.view(I, H), accumulate=True, colsum=gvec[:I]), gemm_fl(I, H))
    case("dgrad ffn-up     [M,3072]x[3072,768] +residual", lambda i: lib.gemm(r(zs, i), W1, b_major=1, residual=r(xs, i), out=out_h), gemm_fl(I, H))
    case("wgrad attn-out   dW[768,768]", lambda i: lib.gemm(r(xs, i), r(xs, i + 1), a_major=1, b_major=1, out=gW[:H * H].view(H, H), accumulate=True), gemm_fl(H, H))
    case("dgrad attn-out   [M,768]x[768,768]", lambda i: lib.gemm(r(xs, i), Wo, b_major=1, out=out_h), gemm_fl(H, H))
    case("attn bwd         (prep + dK/dV + dQ)", lambda i: ops.attention_bwd(qkvs[0], bias, ctx0, r(xs, i), lse0, B, S, NH, pa, seed, 0, dqkv=out_q), 2.5 * attn_f)
    case("attn bwd         bench padding, masked key tiles skipped",
         lambda i: ops.attention_bwd(qkvs[0], bbias, ctx1, r(xs, i), lse1, B, S, NH, pa, seed, 0, dqkv=out_q, key_tiles=plan), 2.5 * attn_f)
    # the step's own layout: packed rows (the padding rows of the bench batch dropped), keep bits stored / read
    pplan = ops.pack_plan(bbias, None, None, B, S)
    n_rows = int(pplan.counts[1])
    lens_p = (pplan.start[1:] - pplan.start[:-1]).double()
    attn_fp = float(4.0 * NH * 64 * (lens_p * lens_p).sum())
    qkv_pk = torch.zeros(pplan.cap, 3 * H, dtype=BF16, device=dev)
    qkv_pk[:n_rows] = qkvs[0][pplan.row_src[:n_rows].long()]
    dctx_pk = torch.zeros(pplan.cap, H, dtype=BF16, device=dev)
    dctx_pk[:n_rows] = xs[0][pplan.row_src[:n_rows].long()]
    ctx_pk, lse_pk = ops.attention_fwd(qkv_pk, None, B, S, NH, pa, seed, 0, plan=pplan, keep_bits=kbits)
    dq_pk = torch.empty_like(qkv_pk)
    case("attn fwd         packed rows (%d of %d), keep bits stored" % (n_rows, M),
         lambda i: ops.attention_fwd(qkv_pk, None, B, S, NH, pa, seed, 0, ctx=ctx_pk, plan=pplan, keep_bits=kbits), attn_fp)
    case("attn bwd         packed rows, keep bits read",
         lambda i: ops.attention_bwd(qkv_pk, None, ctx_pk, dctx_pk, lse_pk, B, S, NH, pa, seed, 0, dqkv=dq_pk, plan=pplan, keep_bits=kbits), 2.5 * attn_fp)
    if kbits is not None:
        case("attn bwd         ... reading the keep bits",
             lambda i: ops.attention_bwd(qkvs[0], bbias, ctx1, r(xs, i), lse1, B, S, NH, pa, seed, 0, dqkv=out_q, key_tiles=plan, keep_bits=kbits), 2.5 * attn_f)
    case("colsum           [M,2304] (separate kernel, no longer in the step)", lambda i: ops.colsum(r(qkvs, i), gvec[:3 * H]), 0, 1.0 * M * 3 * H * 2)
    case("wgrad qkv        dW[2304,768]", lambda i: lib.gemm(r(qkvs, i), r(xs, i), a_major=1, b_major=1, out=gW[:3 * H * H].view(3 * H, H), accumulate=True), gemm_fl(3 * H, H))
    case("wgrad qkv        + bias gradient", lambda i: lib.gemm(r(qkvs, i), r(xs, i), a_major=1, b_major=1, out=gW[:3 * H * H].view(3 * H, H), accumulate=True, colsum=gvec[:3 * H]), gemm_fl(3 * H, H))
    case("dgrad qkv        [M,2304]x[2304,768] +residual", lambda i: lib.gemm(r(qkvs, i), Wqkv, b_major=1, residual=r(xs, i), out=out_h), gemm_fl(3 * H, H))

    # the step before forward (input_pipeline.py), outside the layer total
    from visitron_b200 import input_pipeline as ip
    T, R = 511, 256
    tok = [torch.randint(1000, 30522, (B, T), device=dev) for _ in range(NROT)]
    tcl = torch.full((B, T), -1, dtype=torch.int64, device=dev)
    lut = ip.special_lut(30522, [0, 100, 101, 102, 103], dev)
    cnt = torch.full((B,), 180, dtype=torch.int64)
    offs = torch.cat([torch.zeros(1, dtype=torch.int64), cnt.cumsum(0)]).to(dev)
    rag = [(torch.randn(B * 180, 2054, device=dev), torch.randn(B * 180, 128, device=dev)) for _ in range(NROT)]
    n_layer = len(cases)
    case("mlm mask         [B,511] ids -> ids', labels, attention", lambda i: ip.mask_tokens(r(tok, i), tcl, lut, 103, 0, seed + i), 0, B * T * (8 * 4 + 1.0))
    case("pack regions     180 -> [B,256,2054+128] fp32", lambda i: ip.pack_regions(*r(rag, i), offs, R), 0, 4.0 * B * (180 * 2 + 76) * (2054 + 128))

    # the remaining HBM-bound kernels of the step (embeddings, cross-entropy over the labelled rows, AdamW) and the agents' LSTM
    V, NLAB = 30522, 1408
    word, posw, typw = torch.randn(V, H, device=dev) * 0.02, torch.randn(512, H, device=dev) * 0.02, torch.randn(2, H, device=dev) * 0.02
    gword, gpos, gtyp = torch.zeros_like(word), torch.zeros_like(posw), torch.zeros_like(typw)
    emb_out = torch.empty(B * S, H, dtype=BF16, device=dev)
    case("embed+ln fwd     [B,511] ids -> [B*S,768]", lambda i: ops.embed_ln_fwd(r(tok, i), None, None, word, posw, typw, gamma, beta, 1e-12, emb_out, S, ph, seed, 1000),
         0, B * T * (3 * H * 4 + H * 2.0))
    case("embed+ln bwd     scatter-add into the fp32 tables", lambda i: ops.embed_ln_bwd(r(xs, i), r(tok, i), None, None, word, posw, typw, gamma, 1e-12, S, gword, gpos, gtyp,
                                                                                       gvec[:H], gvec[H:2 * H], ph, seed, 1000), 0, B * T * (H * 2 + 3 * H * 4 + 2 * H * 4.0))
    ldv = ops.roundup(V, 8)
    logits = torch.randn(NLAB, ldv, device=dev)
    clab = torch.randint(0, V, (NLAB,), device=dev)
    g1 = torch.ones(1, device=dev)
    _, _, cstats = ops.softmax_ce_fwd(logits[:, :V], V, clab)
    case("softmax-ce fwd   [1408, 30522] fp32 logits", lambda i: ops.softmax_ce_fwd(logits[:, :V], V, clab), 0, NLAB * V * 4.0)
    case("softmax-ce bwd   -> bf16 dlogits", lambda i: ops.softmax_ce_bwd(logits[:, :V], V, clab, cstats, 1.0 / NLAB, g1, ld_out=ldv), 0, NLAB * V * 6.0)
    from visitron_b200 import optimization as vopt
    pw = torch.nn.Parameter(torch.randn(113_000_000, device=dev) * 0.02)
    pw.grad = torch.randn_like(pw) * 1e-3
    opt = vopt.AdamW([pw], lr=1e-4, weight_decay=0.01)
    case("adamw            113 M fp32 parameters, one launch", lambda i: opt.step(), 0, 28.0 * pw.numel())
    from visitron_b200.agent_models import _LSTMSeqFn
    LB, LT = 8, 511
    lx = rnd(LB * LT, H)
    lw = [torch.randn(2048, 768, device=dev) * 0.03, torch.randn(2048, 512, device=dev) * 0.03, torch.zeros(2048, device=dev), torch.zeros(2048, device=dev)]
    for t_ in lw:
        t_.requires_grad_(True)
    llen = torch.full((LB,), LT, dtype=torch.int64, device=dev)

    def lstm_fwd_bwd(i):
        xi = lx.detach().requires_grad_(True)
        c_, h_, cc_ = _LSTMSeqFn.apply(xi, llen, LB, LT, *lw)
        (c_.sum() + h_.sum()).backward()
    case("lstm 768->512    b=8, 511 steps, fwd + bwd (+ 4 GEMMs)", lstm_fwd_bwd, 3 * LB * LT * 2.0 * (768 + 512) * 2048)

    rows = []
    tot = 0.0
    if args.once:  # for `ncu --profile-from-start off`: one untimed pass, then one profiled launch of every case
        sel = [c for c in cases if not args.only or args.only in c[0]]
        for name, fn, fl, by in sel:
            fn(0)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for name, fn, fl, by in sel:
            fn(1)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    for ci, (name, fn, fl, by) in enumerate(cases):
        if args.only and args.only not in name:
            continue
        if ci == n_layer:
            print("layer total (fwd+bwd, attention with the key-tile plan) %.1f us  -> x12 = %.2f ms" % (tot, tot * 12 / 1e3))
            layer_tot = tot
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / args.iters
        # the layer total counts what the step launches: attention with the key-tile plan, the bias gradients inside the wgrad GEMMs
        skip = (name.startswith("attn") and "packed rows" not in name) or "no longer in the step" in name or \
            name.startswith("wgrad ffn-up     dW") or name.startswith("wgrad qkv        dW") or "dgamma/dbeta/dbias" in name
        tot += us if ci < n_layer and not skip else 0.0
        rate = "%7.0f TFLOP/s" % (fl / us / 1e6) if fl else "%7.0f GB/s   " % (by / us / 1e3)
        print("%-62s %8.1f us  %s" % (name, us, rate), flush=True)
        rows.append(dict(name=name.split("  ")[0].strip(), us=round(us, 1), tflops=round(fl / us / 1e6, 1) if fl else None,
                         gbs=round(by / us / 1e3, 1) if by else None))
    if args.json:
        json.dump(dict(batch=B, seq=S, rows=rows, layer_total_us=tot), open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
