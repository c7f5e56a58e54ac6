"""Diagnostic: packed vs dense attention (forward / backward), where do they differ?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from visitron_b200 import ops
DEV, BF = "cuda", torch.bfloat16
for (B, S, heads, p) in [(2, 64, 1, 0.0), (5, 300, 3, 0.0), (4, 767, 2, 0.1)]:
    torch.manual_seed(B + S)
    Hh = heads * 64
    g = torch.Generator().manual_seed(5)
    lens = torch.randint(1, S + 1, (B,), generator=g); lens[0] = S
    mask = (torch.arange(S)[None, :] < lens[:, None]).long()
    bias = ops.mask_bias(mask.to(DEV))
    qkv = torch.randn(B * S, 3 * Hh, device=DEV).to(BF)
    dctx = torch.randn(B * S, Hh, device=DEV).to(BF)
    ctx, lse = ops.attention_fwd(qkv, bias, B, S, heads, p, 77, 3, key_tiles=ops.attention_key_tiles(bias))
    ctx_nk, _ = ops.attention_fwd(qkv, bias, B, S, heads, p, 77, 3)
    dqkv = ops.attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, p, 77, 3)
    pl = ops.pack_plan(bias, None, None, B, S)
    n = int(pl.counts[1]); src = pl.row_src[:n].long()
    qkv_p = torch.zeros(pl.cap, 3 * Hh, device=DEV, dtype=BF); dctx_p = torch.zeros(pl.cap, Hh, device=DEV, dtype=BF)
    qkv_p[:n], dctx_p[:n] = qkv[src], dctx[src]
    ctx_p, lse_p = ops.attention_fwd(qkv_p, None, B, S, heads, p, 77, 3, plan=pl)
    dqkv_p = ops.attention_bwd(qkv_p, None, ctx_p, dctx_p, lse_p, B, S, heads, p, 77, 3, plan=pl)
    torch.cuda.synchronize()
    print("case", B, S, heads, p, "lens", lens.tolist(), "start", pl.start.tolist(), "counts", pl.counts.tolist(), "plan", [hex(int(v) & 0xffffffff) for v in pl.key_tiles.tolist()])
    print("  dense plan vs no-plan ctx maxdiff", float((ctx.float() - ctx_nk.float()).abs().max()))
    dc = (ctx_p[:n].float() - ctx[src].float()).abs()
    print("  ctx maxdiff", float(dc.max()), "nan", int(torch.isnan(ctx_p[:n].float()).sum()), "rows differing", int((dc.max(1).values > 0).sum()), "of", n)
    rb = pl.row_b[:n].long()
    for b in range(B):
        m = rb == b
        if m.any():
            rows = torch.nonzero(dc[m].max(1).values > 0).squeeze(1)
            print("   sample", b, "len", int(lens[b]), "ctx diff rows", rows[:6].tolist(), "...", int(rows.numel()), "max", float(dc[m].max()))
    dl = (lse_p.float() - lse.float()).abs()
    for b in range(B):
        L = int(lens[b]); print("   lse diff sample", b, float(dl[b, :, :L].max()))
    for name, sl in (("dQ", slice(0, Hh)), ("dK", slice(Hh, 2 * Hh)), ("dV", slice(2 * Hh, 3 * Hh))):
        d = (dqkv_p[:n, sl].float() - dqkv[src][:, sl].float())
        print("  ", name, "maxdiff", float(d.abs().max()), "rel", float(d.norm() / dqkv[src][:, sl].float().norm()), "nan", int(torch.isnan(dqkv_p[:n, sl].float()).sum()))
