"""TEST DOUBLE for the C ABI -- NOT PRODUCT, never imported by visitron_b200.

The build container has no GPU, so the ``-m "not gpu"`` suite cannot call libvisitron_b200.so's
compute entry points.  To still check the HOST logic of visitron_b200/modeling.py (operand majors,
gradient-buffer offsets, residual / remap wiring, bucket order, autograd plumbing) on CPU, this
module restates each tensor-level wrapper of ``visitron_b200.ops`` / ``lib.gemm`` with plain torch
on CPU tensors, keeping the storage dtypes (bf16 activations, fp32 accumulators) so that the
rounding points are the kernels' own.  ``install(monkeypatch)`` swaps them in for one test.
Dropout is only emulated for p == 0 (the parity configuration); p > 0 is a GPU test.
"""
import math

import torch
import torch.nn.functional as F

BF16, F32 = torch.bfloat16, torch.float32
H = 768
CALLS = []


def _remap_rows(n, remap, device):
    m = torch.arange(n, device=device)
    rin, rout, off = remap
    if rin and rin > 0:
        return (m // rin) * rout + off + (m % rin)
    return m


# ---- device-resident row counts (ops.compact_labels / ops.row_limit): rows past the padded count are never touched by the
# kernels, so the double fills them with NaN -- any host logic that lets one leak into a result fails the test ----
_LIMIT = None   # (rows padded to 128, rows) while a row_limit is active


def _rows(n):
    return n if _LIMIT is None else min(n, _LIMIT[0])


def _nan_like(shape, dtype):
    return torch.full(shape, float("nan"), dtype=dtype)


def compact_labels(labels_a, classes_a, labels_b, classes_b, capacity, ignore_index=-1):
    CALLS.append("compact_labels")
    outs, counts = [], torch.zeros(8, dtype=torch.int32)
    for w, (lab, classes) in enumerate(((labels_a, classes_a), (labels_b, classes_b))):
        idx = torch.full((capacity,), -12345, dtype=torch.int64)   # garbage past the padded count
        val = torch.full((capacity,), -12345, dtype=torch.int64)
        nz = torch.nonzero(lab != ignore_index).squeeze(1)
        n = min(int(nz.numel()), capacity)
        npad = min(capacity, (n + 127) // 128 * 128)
        idx[:n], val[:n] = nz[:n], lab[nz[:n]]
        idx[n:npad], val[n:npad] = -1, 0
        counts[4 * w], counts[4 * w + 1] = npad, n
        counts[4 * w + 2] = int(((val[:n] < 0) | (val[:n] >= classes)).sum()) + int(nz.numel() > capacity)
        outs += [idx, val]
    return outs[0], outs[1], outs[2], outs[3], counts


class row_limit:
    def __init__(self, counts, which=0):
        self.v = (int(counts[4 * which]), int(counts[4 * which + 1]))

    def __enter__(self):
        global _LIMIT
        _LIMIT = self.v

    def __exit__(self, *exc):
        global _LIMIT
        _LIMIT = None
        return False


def _no_dropout(p):
    assert not p or p <= 0, "the CPU test double only covers dropout p == 0"


def gemm(A, B, *, a_major=0, b_major=0, out=None, out_dtype=BF16, bias=None, act=0, alpha=1.0, alpha_dev=None,
         residual=None, aux=None, aux_out=None, dropout_p=0.0, seed=0, dropout_stream=0, accumulate=False, split_k=0,
         remap=None, remap_index=None, colsum=None, M=None, N=None, K=None):
    CALLS.append("gemm")
    assert A.dtype == BF16 and B.dtype == BF16
    _no_dropout(dropout_p)
    Af = A.float().t() if a_major else A.float()
    Bf = B.float() if b_major else B.float().t()
    k = min(Af.size(1), Bf.size(0)) if K is None else K
    m_full = Af.size(0)
    if _LIMIT is not None:
        if a_major:
            k = min(k, _LIMIT[0])      # weight-gradient form: the contraction runs over the rows
        else:
            Af = Af[: _LIMIT[0]]
    r = alpha * (Af[:, :k] @ Bf[:k])
    if colsum is not None:
        assert a_major and colsum.dtype == F32
        cs = alpha * Af[:, :k].sum(1)
        colsum[: cs.numel()] += cs[: colsum.numel()] if M is None else cs[:M]
    if M is not None:
        r = r[:M]
    if N is not None:
        r = r[:, :N]
    if alpha_dev is not None:
        r = r * alpha_dev.float().reshape(())
    n = r.size(1)
    if bias is not None:
        assert bias.dtype == F32
        r = r + bias[:n]
    _dgelu = lambda x: 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0))) + x * torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)  # noqa: E731
    if aux_out is not None:
        aux_out[: r.size(0), :n].copy_((_dgelu(r) if act == 5 else r).to(BF16))
    if act in (1, 5):
        r = r * 0.5 * (1.0 + torch.erf(r / math.sqrt(2.0)))
    elif act == 2:
        r = torch.tanh(r)
    elif act == 3:
        x = aux.float()[: r.size(0), :n]
        r = r * (0.5 * (1.0 + torch.erf(x / math.sqrt(2.0))) + x * torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi))
    elif act == 4:
        y = aux.float()[: r.size(0), :n]
        r = r * (1.0 - y * y)
    elif act == 6:
        r = r * aux.float()[: r.size(0), :n]
    rows = _remap_rows(r.size(0), remap or (0, 0, 0), r.device)
    if remap_index is not None:          # packed rows: look the row up, drop it when the map says -1
        rows = remap_index.long()[rows]
        live = rows >= 0
        rows, r = rows[live], r[live]
    if residual is not None:
        r = r + residual[rows].float()[:, :n]
    if out is None:
        assert remap is None and remap_index is None
        if r.size(0) < m_full and not a_major:
            full = _nan_like((m_full, r.size(1)), out_dtype)
            full[: r.size(0)] = r.to(out_dtype)
            return full
        return r.to(out_dtype)
    if accumulate:
        assert out.dtype == F32
        out[rows] = out[rows] + r
    else:
        out[rows] = r.to(out.dtype)
    return out


def _ln(x, g, b, eps):
    u = x.mean(-1, keepdim=True)
    s = (x - u).pow(2).mean(-1, keepdim=True)
    return g * ((x - u) / torch.sqrt(s + eps)) + b


def layernorm_fwd(x, gamma, beta, eps, out=None, dropout_p=0.0, seed=0, stream_id=0, remap=(0, 0, 0)):
    CALLS.append("layernorm_fwd")
    _no_dropout(dropout_p)
    n = _rows(x.size(0))
    y = _ln(x[:n].float(), gamma, beta, eps).to(BF16)
    if out is None:
        if n < x.size(0):
            full = _nan_like(tuple(x.shape), BF16)
            full[:n] = y
            return full
        return y
    out[_remap_rows(n, remap, x.device)] = y
    return out


def layernorm_bwd(dy, x, gamma, eps, dgamma, dbeta, dbias=None, want_dx=True, out_dropout_p=0.0, out_stream=0,
                  in_dropout_p=0.0, in_stream=0, seed=0, remap=(0, 0, 0), dx=None, dxd=None):
    CALLS.append("layernorm_bwd")
    _no_dropout(out_dropout_p)
    _no_dropout(in_dropout_p)
    n = _rows(x.size(0))
    rows = _remap_rows(n, remap, x.device)
    d = dy[rows].float()
    xr = x[:n].float().requires_grad_(True)
    g = gamma.detach().clone().requires_grad_(True)
    b = torch.zeros_like(g).requires_grad_(True)
    with torch.enable_grad():
        _ln(xr, g, b, eps).backward(d)
    dgamma += g.grad
    dbeta += b.grad
    res = xr.grad.to(BF16)
    if dbias is not None:
        dbias += xr.grad.sum(0)
    if n < x.size(0):
        full = _nan_like(tuple(x.shape), BF16)
        full[:n] = res
        res = full
    return res, res


def _embed_rows(b, T, rows_per_sample, row_dst):
    """flat output row of every (sample, position) and which of them exist"""
    o = (torch.arange(b)[:, None] * rows_per_sample + torch.arange(T)[None, :]).reshape(-1)
    if row_dst is None:
        return o, torch.ones_like(o, dtype=torch.bool)
    o = row_dst.long()[o]
    return o.clamp_min(0), o >= 0


def embed_ln_fwd(ids, type_ids, pos_ids, word, pos, typ, gamma, beta, eps, out, rows_per_sample, dropout_p=0.0, seed=0, stream_id=0, row_dst=None):
    CALLS.append("embed_ln_fwd")
    _no_dropout(dropout_p)
    b, T = ids.shape
    p = torch.arange(T)[None].expand(b, T) if pos_ids is None else pos_ids
    e = word[ids] + pos[p] + (typ[type_ids] if type_ids is not None else typ[0])
    o, live = _embed_rows(b, T, rows_per_sample, row_dst)
    out[o[live]] = _ln(e, gamma, beta, eps).to(BF16).reshape(b * T, H)[live]
    return out


def embed_ln_bwd(dy, ids, type_ids, pos_ids, word, pos, typ, gamma, eps, rows_per_sample, dword, dpos, dtype, dgamma, dbeta,
                 dropout_p=0.0, seed=0, stream_id=0, row_dst=None):
    CALLS.append("embed_ln_bwd")
    _no_dropout(dropout_p)
    b, T = ids.shape
    p = torch.arange(T)[None].expand(b, T) if pos_ids is None else pos_ids
    tt = torch.zeros_like(ids) if type_ids is None else type_ids
    ws = [t.detach().clone().requires_grad_(True) for t in (word, pos, typ, gamma)]
    beta = torch.zeros_like(gamma).requires_grad_(True)
    o, live = _embed_rows(b, T, rows_per_sample, row_dst)
    g = torch.where(live[:, None], dy[o].float(), torch.zeros(()))       # a dropped row has no gradient
    with torch.enable_grad():
        y = _ln(ws[0][ids] + ws[1][p] + ws[2][tt], ws[3], beta, eps)
        y.backward(g.view(b, T, H))
    dword += ws[0].grad
    dpos += ws[1].grad
    dtype += ws[2].grad
    dgamma += ws[3].grad
    dbeta += beta.grad


def colsum(x, out, ncols=None, accumulate=True):
    CALLS.append("colsum")
    n = x.size(1) if ncols is None else ncols
    s = x.float()[: _rows(x.size(0)), :n].sum(0)
    if accumulate:
        out[:n] += s
    else:
        out[:n] = s
    return out


def cast_weight(src, dst=None, dst_t=None):
    CALLS.append("cast_weight")
    if dst is not None:
        dst[:, : src.size(1)] = src.to(BF16)
    if dst_t is not None:
        dst_t[:, : src.size(0)] = src.t().to(BF16)


def pad_cast_concat(a, b, out):
    CALLS.append("pad_cast_concat")
    out.zero_()
    out[:, : a.size(1)] = a.to(BF16)
    if b is not None:
        out[:, a.size(1): a.size(1) + b.size(1)] = b.to(BF16)
    return out


def dropout_rows(x, rows, out, dropout_p, seed, stream_id, remap=(0, 0, 0), row_dst=None):
    CALLS.append("dropout_rows")
    _no_dropout(dropout_p)
    src = _remap_rows(rows, remap, x.device)
    if row_dst is None:
        out[:rows] = x[src]
    else:
        src = row_dst.long()[src]
        out[:rows] = torch.where((src >= 0)[:, None], x[src.clamp_min(0)], torch.zeros((), dtype=x.dtype))
    return out


def gather_rows(src, idx, out=None):
    CALLS.append("gather_rows")
    n = _rows(idx.numel())
    r = _nan_like((idx.numel(), src.size(1)), src.dtype)
    live = idx[:n]
    r[:n] = torch.where((live >= 0)[:, None], src[live.clamp_min(0)], torch.zeros((), dtype=src.dtype))
    if out is None:
        return r.contiguous()
    out.copy_(r)
    return out


def scatter_add_rows(src, idx, dst):
    CALLS.append("scatter_add_rows")
    n = _rows(idx.numel())
    keep = idx[:n] >= 0
    ii = idx[:n][keep]
    dst[ii] = (dst[ii].float() + src[:n][keep].float()).to(BF16)


def mask_bias(mask):
    CALLS.append("mask_bias")
    return (1.0 - mask.to(F32)) * -10000.0


def _attn(qkv, bias, B, S, heads):
    Hh = heads * 64
    q, k, v = [t.view(B, S, heads, 64).permute(0, 2, 1, 3) for t in qkv.view(B * S, 3, Hh).unbind(1)]
    s = q @ k.transpose(-1, -2) / 8.0 + bias[:, None, None, :]
    p = torch.softmax(s, -1)
    return (p @ v).permute(0, 2, 1, 3).reshape(B * S, Hh)


def attention_key_tiles(bias):
    CALLS.append("attention_key_tiles")
    return None


class PackPlan:
    pass


def pack_plan(bias, labels_a, labels_b, B, S):
    """CPU restatement of vt_pack_plan (csrc/pack.cu): keep a position iff mask != 0, labelled, or position 0."""
    CALLS.append("pack_plan")
    pl = PackPlan()
    M = B * S
    keep = bias.reshape(-1) != -10000.0
    for lab in (labels_a, labels_b):
        if lab is not None:
            keep = keep | (lab.reshape(-1) != -1)
    keep = keep | (torch.arange(M) % S == 0)
    pl.B, pl.S = B, S
    pl.cap = cap = (M + 127) // 128 * 128 + 128
    src = torch.nonzero(keep).squeeze(1)
    n = src.numel()
    pl.row_src = torch.full((cap,), -1, dtype=torch.int32)
    pl.row_src[:n] = src.int()
    pl.row_dst = torch.full((M,), -1, dtype=torch.int32)
    pl.row_dst[src] = torch.arange(n, dtype=torch.int32)
    pl.row_b = torch.full((cap,), -1, dtype=torch.int32)
    pl.row_b[:n] = (src // S).int()
    per = keep.view(B, S).sum(1)
    pl.start = torch.cat([torch.zeros(1, dtype=torch.long), per.cumsum(0)]).int()
    pl.counts = torch.tensor([min(cap, (n + 127) // 128 * 128 + 128), n, 0, 0], dtype=torch.int32)
    pl.bias = torch.zeros(cap)
    pl.bias[:n] = bias.reshape(-1)[src]
    pl.labels_a = pl.labels_b = None
    if labels_a is not None:
        pl.labels_a = torch.full((cap,), -1, dtype=torch.int64)
        pl.labels_a[:n] = labels_a.reshape(-1)[src]
    if labels_b is not None:
        pl.labels_b = torch.full((cap,), -1, dtype=torch.int64)
        pl.labels_b[:n] = labels_b.reshape(-1)[src]
    pl.cls_rows = pl.start[:B].long()
    pl.key_tiles = None
    return pl


def zero_pad_rows(buf, counts):
    CALLS.append("zero_pad_rows")
    buf[int(counts[1]): int(counts[0])] = 0
    return buf


def _attn_packed(qkv, plan, heads):
    """per-sample attention over the packed rows; rows that belong to no sample stay NaN (the kernels never write them)"""
    out = _nan_like((qkv.size(0), heads * 64), F32)
    parts = []
    for b in range(plan.B):
        s0, s1 = int(plan.start[b]), int(plan.start[b + 1])
        parts.append((s0, s1, _attn(qkv[s0:s1], plan.bias[None, s0:s1], 1, s1 - s0, heads)))
    return out, parts


def attention_fwd(qkv, bias, B, S, heads, dropout_p=0.0, seed=0, stream_id=0, ctx=None, key_tiles=None, plan=None, keep_bits=None):
    CALLS.append("attention_fwd")
    _no_dropout(dropout_p)
    if plan is None:
        out = _attn(qkv.float(), bias, B, S, heads).to(BF16)
        return out, torch.zeros((B, heads, S, 2))
    out, parts = _attn_packed(qkv.float(), plan, heads)
    for s0, s1, o in parts:
        out[s0:s1] = o
    out = out.to(BF16)
    zero_pad_rows(out, plan.counts)
    return out, torch.zeros((B, heads, S, 2))


def attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=0.0, seed=0, stream_id=0, dqkv=None, key_tiles=None, plan=None, keep_bits=None):
    CALLS.append("attention_bwd")
    _no_dropout(dropout_p)
    q = qkv.float().requires_grad_(True)
    if plan is None:
        with torch.enable_grad():
            _attn(q, bias, B, S, heads).backward(dctx.float())
        return q.grad.to(BF16)
    with torch.enable_grad():
        _, parts = _attn_packed(q, plan, heads)
        sum((o * dctx[s0:s1].float()).sum() for s0, s1, o in parts).backward()
    d = _nan_like(tuple(qkv.shape), BF16)
    n = int(plan.counts[1])
    d[:n] = q.grad[:n].to(BF16)
    zero_pad_rows(d, plan.counts)
    return d


def softmax_ce_fwd(logits, ncols, labels, double_softmax=False):
    CALLS.append("softmax_ce_fwd")
    rows = logits.size(0)
    n = rows if _LIMIT is None else min(rows, _LIMIT[1])
    x = logits[:n, :ncols].float()
    inp = torch.softmax(x, -1) if double_softmax else x
    loss = _nan_like((rows,), F32)
    arg = torch.full((rows,), -12345, dtype=torch.int64)
    if n:
        loss[:n] = F.cross_entropy(inp, labels[:n], reduction="none")
        arg[:n] = x.argmax(1)
    return loss, arg, torch.zeros((rows, 2))


def softmax_ce_bwd(logits, ncols, labels, stats, scale, scale_dev, double_softmax=False, ld_out=None):
    CALLS.append("softmax_ce_bwd")
    rows = logits.size(0)
    n, npad = (rows, rows) if _LIMIT is None else (min(rows, _LIMIT[1]), min(rows, _LIMIT[0]))
    if _LIMIT is not None:
        scale = scale / max(n, 1)      # the mean's 1 / n is applied on the device
    x = logits[:n, :ncols].float().detach().clone().requires_grad_(True)
    ld_out = (ncols + 7) // 8 * 8 if ld_out is None else ld_out
    d = _nan_like((rows, ld_out), BF16)
    d[:npad] = 0
    if n:
        with torch.enable_grad():
            inp = torch.softmax(x, -1) if double_softmax else x
            (F.cross_entropy(inp, labels[:n], reduction="sum") * scale * scale_dev.reshape(())).backward()
        d[:n, :ncols] = x.grad.to(BF16)
    return d


def action_head_fwd(pooled, W, bias, labels):
    CALLS.append("action_head_fwd")
    sc = torch.log_softmax(F.linear(pooled, W, bias), -1)
    if labels is None:
        loss = sc.new_zeros(sc.size(0))
    else:
        safe = labels.clamp(min=0)
        loss = F.cross_entropy(sc, safe, reduction="none") * (labels >= 0)
    return sc, loss, sc.argmax(1)


def action_head_bwd(pooled, W, labels, scores, n_valid_dev, g_dev, dW, dbias):
    CALLS.append("action_head_bwd")
    p = pooled.detach().clone().requires_grad_(True)
    w = W.detach().clone().requires_grad_(True)
    b = torch.zeros(W.size(0)).requires_grad_(True)
    with torch.enable_grad():
        sc = torch.log_softmax(F.linear(p, w, b), -1)
        l = F.cross_entropy(sc, labels, ignore_index=-1, reduction="sum") / n_valid_dev.reshape(()) * g_dev.reshape(())
        l.backward()
    dW += w.grad
    dbias += b.grad
    return p.grad


def pretrain_finalize(mlm, tok, act):
    CALLS.append("pretrain_finalize")
    out = torch.zeros(8)
    (ml, ma, mlab), (tl, ta, tlab), (al, aa, alab) = mlm, tok, act
    out[1] = ml.sum() / ml.numel() if ml.numel() else float("nan")
    out[3] = tl.sum() / tl.numel() if tl.numel() else float("nan")
    out[4] = (ma == mlab).float().sum() / max(ml.numel(), 1) if ml.numel() else float("nan")
    out[6] = (ta == tlab).float().sum() / max(tl.numel(), 1) if tl.numel() else float("nan")
    if al.numel():
        valid = (alab >= 0).float()
        out[7] = valid.sum()
        out[2] = (al * valid).sum() / valid.sum()
        out[5] = (aa == alab).float().sum() / al.numel()
    out[0] = out[1] + out[2] + out[3]
    return out


def pretrain_finalize_counted(mlm, tok, act, counts):
    CALLS.append("pretrain_finalize_counted")
    n, nt = int(counts[1]), int(counts[5])
    out = pretrain_finalize(tuple(t[:n] for t in mlm), tuple(t[:nt] for t in tok), act)
    if int(counts[2]) or int(counts[6]):
        out[0] = float("nan")
    return out


def tanh_bwd(dy, y):
    CALLS.append("tanh_bwd")
    return (dy * (1.0 - y * y)).to(BF16)


def gelu_bwd(dy, pre):
    CALLS.append("gelu_bwd")
    n = _rows(pre.size(0))
    x = pre[:n].float()
    g = 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0))) + x * torch.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)
    out = _nan_like(tuple(pre.shape), BF16)
    out[:n] = (dy[:n].float() * g).to(BF16)
    return out


def to_f32(x):
    CALLS.append("to_f32")
    return x.float()


def to_bf16(x, addend=None, out=None):
    CALLS.append("to_bf16")
    y = x.to(BF16) if addend is None else (x + addend.float()).to(BF16)
    if out is None:
        return y
    out.copy_(y)
    return out


_OPS = ["layernorm_fwd", "layernorm_bwd", "embed_ln_fwd", "embed_ln_bwd", "colsum", "cast_weight", "pad_cast_concat",
        "dropout_rows", "gather_rows", "scatter_add_rows", "mask_bias", "attention_key_tiles", "attention_fwd", "attention_bwd", "softmax_ce_fwd",
        "softmax_ce_bwd", "action_head_fwd", "action_head_bwd", "pretrain_finalize", "pretrain_finalize_counted", "compact_labels", "row_limit",
        "pack_plan", "zero_pad_rows",
        "tanh_bwd", "gelu_bwd", "to_f32", "to_bf16"]


def install(monkeypatch):
    """Swap the C-ABI wrappers for the CPU restatements (one test's lifetime)."""
    from visitron_b200 import lib, ops
    g = globals()
    for name in _OPS:
        assert hasattr(ops, name), name
        monkeypatch.setattr(ops, name, g[name])
    monkeypatch.setattr(lib, "gemm", gemm)
    monkeypatch.setattr(lib, "require_device", lambda t: None)
    del CALLS[:]
