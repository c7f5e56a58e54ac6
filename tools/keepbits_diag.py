"""bits-vs-regenerate backward: where do they differ?"""
import torch
from visitron_b200 import ops
DEV, BF = "cuda", torch.bfloat16
def run(B, S, heads, packed, p=0.1):
    torch.manual_seed(1)
    Hh = heads * 64
    g = torch.Generator().manual_seed(5)
    lens = torch.randint(1, S + 1, (B,), generator=g); lens[0] = S
    if not packed: lens[:] = S
    mask = (torch.arange(S)[None, :] < lens[:, None]).long()
    bias = ops.mask_bias(mask.to(DEV))
    qkv = torch.randn(B * S, 3 * Hh, device=DEV).to(BF)
    dctx = torch.randn(B * S, Hh, device=DEV).to(BF)
    if packed:
        pl = ops.pack_plan(bias, None, None, B, S)
        n = int(pl.counts[1]); src = pl.row_src[:n].long()
        q = torch.zeros(pl.cap, 3 * Hh, device=DEV, dtype=BF); d = torch.zeros(pl.cap, Hh, device=DEV, dtype=BF)
        q[:n], d[:n] = qkv[src], dctx[src]
        kw = dict(plan=pl); bias_a = None
    else:
        q, d, n, kw, bias_a = qkv, dctx, B * S, {}, bias
    bits = ops.attention_keep_bits(B, S, heads, DEV); bits.fill_(-1)
    ctx, lse = ops.attention_fwd(q, bias_a, B, S, heads, p, 77, 3, keep_bits=bits, **kw)
    a = ops.attention_bwd(q, bias_a, ctx, d, lse, B, S, heads, p, 77, 3, keep_bits=bits, **kw)[:n].float()
    b = ops.attention_bwd(q, bias_a, ctx, d, lse, B, S, heads, p, 77, 3, **kw)[:n].float()
    a2 = ops.attention_bwd(q, bias_a, ctx, d, lse, B, S, heads, p, 77, 3, keep_bits=bits, **kw)[:n].float()
    b2 = ops.attention_bwd(q, bias_a, ctx, d, lse, B, S, heads, p, 77, 3, **kw)[:n].float()
    print("   run-to-run: bits %d regen %d differing elements" % (int((a != a2).sum()), int((b != b2).sum())))
    diff = (a != b)
    print("B %d S %d heads %d packed %s lens %s: differing elements %d of %d" % (B, S, heads, packed, lens.tolist(), int(diff.sum()), diff.numel()))
    if diff.any():
        rows = torch.nonzero(diff.any(1)).squeeze(1)
        el = torch.nonzero(diff)
        for r_, c_ in el[:4].tolist():
            print("   [%d, %d] bits %.9g regen %.9g" % (r_, c_, float(a[r_, c_]), float(b[r_, c_])))
        print("  rows", rows[:10].tolist(), "...", rows[-5:].tolist(), "count", rows.numel())
        for part, nm in enumerate("qkv"):
            dd = diff[:, part * Hh:(part + 1) * Hh]
            print("  d%s: %d elements, max abs diff %.4g" % (nm, int(dd.sum()), float((a - b)[:, part * Hh:(part + 1) * Hh].abs().max())))
for cfg in [(2, 128, 2, False), (2, 256, 1, False), (2, 767, 2, False), (1, 767, 1, True), (4, 767, 2, True), (3, 300, 1, True)]:
    run(*cfg)

def mask_check(B=2, S=767, heads=2, p=0.1, blk=10):
    torch.manual_seed(1)
    Hh = heads * 64
    bias = torch.zeros(B, S, device=DEV)
    qkv = torch.randn(B * S, 3 * Hh, device=DEV).to(BF)
    bits = ops.attention_keep_bits(B, S, heads, DEV); bits.fill_(-1)
    ops.attention_fwd(qkv, bias, B, S, heads, p, 77, 3, keep_bits=bits)
    probe = qkv.clone().view(B, S, 3, heads, 64)
    probe[:, :, 2] = 0
    k0 = blk * 64
    for j in range(min(64, S - k0)):
        probe[:, k0 + j, 2, :, j] = 1
    o, _ = ops.attention_fwd(probe.view(B * S, 3 * Hh), bias, B, S, heads, p, 77, 3)
    rec = (o.view(B, S, heads, 64).permute(0, 2, 1, 3) != 0)          # [B, heads, S(q), 64]
    G = 4 * ((S + 127) // 128)
    w = bits.view(B * heads, G, S).long() & 0xffffffff
    un = ((w[:, :, :, None] >> torch.arange(32, device=DEV)) & 1).permute(0, 2, 1, 3).reshape(B, heads, S, G * 32)[..., k0:k0 + 64].bool()
    n = min(64, S - k0)
    mism = torch.nonzero(un[..., :n] != rec[..., :n])
    print("bits vs recovered mask, keys %d..%d: %d mismatches" % (k0, k0 + n - 1, mism.size(0)), mism[:8].tolist())
    o2 = o.view(B, S, heads, 64).permute(0, 2, 1, 3)
    for m in mism[:4].tolist():
        print("   probability there:", float(o2[m[0], m[1], m[2], m[3]]), "bit", bool(un[m[0], m[1], m[2], m[3]]))
mask_check()
