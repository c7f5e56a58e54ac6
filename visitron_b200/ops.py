"""Tensor-level wrappers over the C ABI (one function per entry point of include/visitron_b200.h).

These allocate outputs with torch and pass raw pointers + the current stream; no arithmetic happens
in Python.  `lib.gemm` lives in lib.py.
"""
from __future__ import annotations

import os

from typing import Optional, Tuple

import torch

from . import lib
from .lib import check, init, ptr, stream_ptr

H = 768
BF16, F32 = torch.bfloat16, torch.float32

_MASK_CODES = {torch.int64: 0, torch.uint8: 1, torch.bool: 1, torch.int32: 2, torch.float32: 3}


def _ws(device, n) -> torch.Tensor:
    return torch.empty(n, device=device, dtype=F32)


def roundup(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def layernorm_fwd(x, gamma, beta, eps, out=None, dropout_p=0.0, seed=0, stream_id=0, remap=(0, 0, 0)):
    L = init()
    rows = x.size(0)
    if out is None:
        out = torch.empty_like(x)
    check(L.vt_layernorm_fwd(ptr(x), x.stride(0), ptr(gamma), ptr(beta), eps, rows, x.size(1), ptr(out), out.stride(0),
                             dropout_p, seed, stream_id, *remap, stream_ptr()), "vt_layernorm_fwd")
    return out


def layernorm_bwd(dy, x, gamma, eps, dgamma, dbeta, dbias=None, want_dx=True, out_dropout_p=0.0, out_stream=0,
                  in_dropout_p=0.0, in_stream=0, seed=0, remap=(0, 0, 0), dx=None, dxd=None):
    """Returns (dx, dx_dropped); dx_dropped is dx itself when no output dropout is active."""
    L = init()
    rows = x.size(0)
    if want_dx and dx is None:
        dx = torch.empty_like(x)
    if out_dropout_p > 0 and dxd is None:
        dxd = torch.empty_like(x)
    parts = _ws(x.device, int(L.vt_row_partials_rows(rows)) * 4 * H)
    check(L.vt_layernorm_bwd(ptr(dy), dy.stride(0), ptr(x), x.stride(0), ptr(gamma), eps, rows, x.size(1),
                             ptr(dx), dx.stride(0) if dx is not None else 0, ptr(dxd), dxd.stride(0) if dxd is not None else 0,
                             ptr(parts), ptr(dgamma), ptr(dbeta), ptr(dbias), in_dropout_p, in_stream, out_dropout_p, out_stream,
                             seed, *remap, stream_ptr()), "vt_layernorm_bwd")
    return dx, (dxd if dxd is not None else dx)


def embed_ln_fwd(ids, type_ids, pos_ids, word, pos, typ, gamma, beta, eps, out, rows_per_sample, dropout_p=0.0, seed=0, stream_id=0, row_dst=None):
    L = init()
    b, T = ids.shape
    if row_dst is None:
        check(L.vt_embed_ln_fwd(ptr(ids), ptr(type_ids), ptr(pos_ids), ptr(word), ptr(pos), ptr(typ), ptr(gamma), ptr(beta), eps,
                                b, T, rows_per_sample, word.size(1), ptr(out), dropout_p, seed, stream_id, stream_ptr()), "vt_embed_ln_fwd")
    else:
        check(L.vt_embed_ln_fwd_packed(ptr(ids), ptr(type_ids), ptr(pos_ids), ptr(word), ptr(pos), ptr(typ), ptr(gamma), ptr(beta), eps,
                                       b, T, rows_per_sample, word.size(1), ptr(out), dropout_p, seed, stream_id, ptr(row_dst), stream_ptr()),
              "vt_embed_ln_fwd_packed")
    return out


def embed_ln_bwd(dy, ids, type_ids, pos_ids, word, pos, typ, gamma, eps, rows_per_sample, dword, dpos, dtype, dgamma, dbeta,
                 dropout_p=0.0, seed=0, stream_id=0, row_dst=None):
    L = init()
    b, T = ids.shape
    parts = _ws(dy.device, int(L.vt_row_partials_rows(b * T)) * 4 * H)
    args = (ptr(dy), ptr(ids), ptr(type_ids), ptr(pos_ids), ptr(word), ptr(pos), ptr(typ), ptr(gamma), eps, b, T,
            rows_per_sample, word.size(1), typ.size(0), ptr(dword), ptr(dpos), ptr(dtype), ptr(dgamma), ptr(dbeta),
            ptr(parts), dropout_p, seed, stream_id)
    if row_dst is None:
        check(L.vt_embed_ln_bwd(*args, stream_ptr()), "vt_embed_ln_bwd")
    else:
        check(L.vt_embed_ln_bwd_packed(*args, ptr(row_dst), stream_ptr()), "vt_embed_ln_bwd_packed")


def colsum(x, out, ncols=None, accumulate=True):
    L = init()
    rows = x.size(0)
    ncols = x.size(1) if ncols is None else ncols
    parts = _ws(x.device, int(L.vt_colsum_partials_rows(rows)) * roundup(ncols, 8))
    check(L.vt_colsum_bf16(ptr(x), x.stride(0), rows, ncols, ptr(parts), ptr(out), 1 if accumulate else 0, stream_ptr()), "vt_colsum_bf16")
    return out


def cast_weight(src, dst=None, dst_t=None):
    """fp32 [rows, cols] -> bf16 dst [rows, >=cols] and/or transposed dst_t [cols, >=rows]."""
    L = init()
    rows, cols = src.shape
    check(L.vt_cast_f32_bf16(ptr(src), src.stride(0), rows, cols, ptr(dst), dst.stride(0) if dst is not None else 0,
                             ptr(dst_t), dst_t.stride(0) if dst_t is not None else 0, stream_ptr()), "vt_cast_f32_bf16")


def pad_cast_concat(a, b, out):
    """out[:, :d0] = a, out[:, d0:d0+d1] = b, zero padding behind; a / b fp32 or bf16 (a bf16 feature cache is read as it is)."""
    L = init()
    rows = a.size(0)
    for t in (a, b):
        if t is not None and t.dtype not in (F32, BF16):
            raise lib.VtError("pad_cast_concat takes fp32 or bf16 sources, got %s" % t.dtype)
    check(L.vt_pad_cast_concat_mixed(ptr(a), int(a.dtype == BF16), a.size(1), ptr(b), int(b is not None and b.dtype == BF16),
                                     b.size(1) if b is not None else 0, rows, ptr(out), out.stride(0), stream_ptr()), "vt_pad_cast_concat")
    return out


def dropout_rows(x, rows, out, dropout_p, seed, stream_id, remap=(0, 0, 0), row_dst=None):
    L = init()
    if row_dst is None:
        check(L.vt_dropout_rows(ptr(x), x.stride(0), rows, H, ptr(out), out.stride(0), dropout_p, seed, stream_id, *remap, stream_ptr()),
              "vt_dropout_rows")
    else:
        check(L.vt_dropout_rows_packed(ptr(x), x.stride(0), rows, H, ptr(out), out.stride(0), dropout_p, seed, stream_id, *remap, ptr(row_dst),
                                       stream_ptr()), "vt_dropout_rows_packed")
    return out


def gather_rows(src, idx, out=None):
    L = init()
    n = idx.numel()
    if out is None:
        out = torch.empty((n, H), device=src.device, dtype=BF16)
    check(L.vt_gather_rows(ptr(src), src.stride(0), ptr(idx), n, H, ptr(out), out.stride(0), stream_ptr()), "vt_gather_rows")
    return out


def scatter_add_rows(src, idx, dst):
    L = init()
    check(L.vt_scatter_add_rows(ptr(src), src.stride(0), ptr(idx), idx.numel(), H, ptr(dst), dst.stride(0), stream_ptr()), "vt_scatter_add_rows")


def mask_bias(mask: torch.Tensor) -> torch.Tensor:
    L = init()
    if mask.dtype not in _MASK_CODES:
        mask = mask.to(torch.float32)
    mask = mask.contiguous()
    out = torch.empty(mask.shape, device=mask.device, dtype=F32)
    check(L.vt_mask_bias(ptr(mask), _MASK_CODES[mask.dtype], mask.numel(), ptr(out), stream_ptr()), "vt_mask_bias")
    return out


def attention_key_tiles(bias: torch.Tensor) -> Optional[torch.Tensor]:
    """uint32-as-int32 [B]: which 64-key tiles of every sample hold an unmasked key; the attention kernels skip the others (exactly-zero
    probabilities in the reference's fp32 softmax, so no bit of the result changes).  None for S > 2048 (every tile is visited)."""
    B, S = bias.shape
    if S > 2048:
        return None
    plan = torch.empty(B, device=bias.device, dtype=torch.int32)
    check(init().vt_attention_key_tiles(ptr(bias), B, S, ptr(plan), stream_ptr()), "vt_attention_key_tiles")
    return plan


class PackPlan:
    """Device-resident description of the packed row layout of one step (vt_pack_plan): no host sync, fixed shapes."""
    __slots__ = ("B", "S", "cap", "row_src", "row_dst", "row_b", "start", "counts", "bias", "labels_a", "labels_b", "cls_rows", "key_tiles", "bwd_sched")


def pack_plan(bias, labels_a, labels_b, B, S) -> PackPlan:
    """Which rows of the padded [B, S] layout exist (mask != 0, labelled, or position 0) and where they live in the packed layout."""
    L = init()
    dev = bias.device
    pl = PackPlan()
    pl.B, pl.S = B, S
    pl.cap = cap = roundup(B * S, 128) + 128
    i32 = torch.empty(2 * cap + B * S + (B + 1) + 4, device=dev, dtype=torch.int32)
    pl.row_src, pl.row_b = i32[:cap], i32[cap:2 * cap]
    pl.row_dst = i32[2 * cap: 2 * cap + B * S]
    pl.start = i32[2 * cap + B * S: 2 * cap + B * S + B + 1]
    pl.counts = torch.empty(8, device=dev, dtype=torch.int32)[:4]       # own 16-byte aligned allocation: it is a row-limit block
    pl.bias = torch.empty(cap, device=dev, dtype=F32)
    i64 = torch.empty(2 * cap + B, device=dev, dtype=torch.int64)
    pl.labels_a = i64[:cap] if labels_a is not None else None
    pl.labels_b = i64[cap:2 * cap] if labels_b is not None else None
    pl.cls_rows = i64[2 * cap:]
    check(L.vt_pack_plan(ptr(bias), ptr(labels_a), ptr(labels_b), B, S, cap, ptr(pl.row_src), ptr(pl.row_dst), ptr(pl.row_b), ptr(pl.start),
                         ptr(pl.counts), ptr(pl.bias), ptr(pl.labels_a), ptr(pl.labels_b), ptr(pl.cls_rows), stream_ptr()), "vt_pack_plan")
    pl.bwd_sched = {}      # heads -> (int32 [ctas + 1], ctas): cost-balanced item ranges of the attention backward, built on first use
    pl.key_tiles = None
    if S <= 2048:
        pl.key_tiles = torch.empty(B, device=dev, dtype=torch.int32)
        check(L.vt_attention_key_tiles_packed(ptr(pl.bias), ptr(pl.start), B, S, ptr(pl.key_tiles), stream_ptr()), "vt_attention_key_tiles_packed")
    return pl


def zero_pad_rows(buf, counts):
    """rows [counts[1], counts[0]) of a 2-D row-major buffer := 0 (the packed layout's padding rows)."""
    check(init().vt_zero_pad_rows(ptr(buf), buf.stride(0) * buf.element_size(), buf.size(1) * buf.element_size(), ptr(counts), stream_ptr()),
          "vt_zero_pad_rows")
    return buf


_NO_BWD_SCHED = bool(int(os.environ.get("VT_NO_BWD_SCHED", "0")))    # A/B switch: round-robin items in the packed attention backward


def attention_keep_bits(B, S, heads, device):
    """Store for the keep decisions of one attention call's probability dropout (see vt_set_attention_keep_bits): pass it to
    attention_fwd (which fills it) and to the matching attention_bwd (which then skips the generator)."""
    return torch.empty(int(init().vt_attention_keep_bits_words(B, S, heads)), device=device, dtype=torch.int32)


class _keep_bits:
    def __init__(self, bits):
        self.bits = bits

    def __enter__(self):
        if self.bits is not None:
            init().vt_set_attention_keep_bits(self.bits.data_ptr())

    def __exit__(self, *exc):
        if self.bits is not None:
            init().vt_set_attention_keep_bits(None)
        return False


def attention_fwd(qkv, bias, B, S, heads, dropout_p=0.0, seed=0, stream_id=0, ctx=None, key_tiles=None, plan: Optional[PackPlan] = None,
                  keep_bits=None):
    L = init()
    if keep_bits is not None and keep_bits.numel() < int(L.vt_attention_keep_bits_words(B, S, heads)):
        raise lib.VtError("keep_bits: too small, see attention_keep_bits")
    rows = B * S if plan is None else plan.cap
    if ctx is None:
        ctx = torch.empty((rows, heads * 64), device=qkv.device, dtype=BF16)
    lse = torch.empty((B, heads, S, 2), device=qkv.device, dtype=F32)
    with _keep_bits(keep_bits):
        if plan is None:
            check(L.vt_attention_fwd(ptr(qkv), qkv.stride(0), ptr(bias), ptr(key_tiles), B, S, heads, 64, ptr(ctx), ptr(lse), dropout_p, seed, stream_id,
                                     stream_ptr()), "vt_attention_fwd")
        else:
            check(L.vt_attention_fwd_packed(ptr(qkv), qkv.stride(0), ptr(plan.bias), ptr(plan.key_tiles), ptr(plan.start), plan.cap, B, S, heads, 64,
                                            ptr(ctx), ptr(lse), dropout_p, seed, stream_id, stream_ptr()), "vt_attention_fwd_packed")
    if plan is not None:
        zero_pad_rows(ctx, plan.counts)
    return ctx, lse


def attention_bwd(qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p=0.0, seed=0, stream_id=0, dqkv=None, key_tiles=None,
                  plan: Optional[PackPlan] = None, keep_bits=None):
    L = init()
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    sched = None
    if plan is not None and B <= 4000 and not _NO_BWD_SCHED:
        sched = plan.bwd_sched.get(heads)
        if sched is None:
            ctas = int(L.vt_attention_bwd_schedule_ctas(B, S, heads))
            buf = torch.empty(ctas + 1, device=qkv.device, dtype=torch.int32)
            check(L.vt_attention_bwd_schedule(ptr(plan.start), B, S, heads, ptr(buf), stream_ptr()), "vt_attention_bwd_schedule")
            sched = plan.bwd_sched[heads] = (buf, ctas)
        L.vt_set_attention_bwd_schedule(sched[0].data_ptr(), sched[1])
    try:
        dqkv = _attention_bwd_launch(L, qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p, seed, stream_id, dqkv, key_tiles, plan, keep_bits)
    finally:
        if sched is not None:
            L.vt_set_attention_bwd_schedule(None, 0)
    return dqkv


def _attention_bwd_launch(L, qkv, bias, ctx, dctx, lse, B, S, heads, dropout_p, seed, stream_id, dqkv, key_tiles, plan, keep_bits):
    with _keep_bits(keep_bits):
        if plan is None:
            ws = torch.empty(int(L.vt_attention_bwd_workspace(B, S, heads)), device=qkv.device, dtype=F32)
            check(L.vt_attention_bwd(ptr(qkv), qkv.stride(0), ptr(bias), ptr(key_tiles), ptr(ctx), ptr(dctx), ptr(lse), ptr(ws), B, S, heads, 64, ptr(dqkv),
                                     dropout_p, seed, stream_id, stream_ptr()), "vt_attention_bwd")
        else:
            ws = torch.empty(int(L.vt_attention_bwd_workspace_packed(plan.cap, B, S, heads)), device=qkv.device, dtype=F32)
            check(L.vt_attention_bwd_packed(ptr(qkv), qkv.stride(0), ptr(plan.bias), ptr(plan.key_tiles), ptr(plan.start), ptr(plan.row_b), plan.cap,
                                            ptr(ctx), ptr(dctx), ptr(lse), ptr(ws), B, S, heads, 64, ptr(dqkv), dropout_p, seed, stream_id, stream_ptr()),
                  "vt_attention_bwd_packed")
    if plan is not None:
        zero_pad_rows(dqkv, plan.counts)
    return dqkv


# ------------------------------------------------------------------------------------------------
# device-resident step state (sync-free steps, CUDA graphs): see the first section of include/visitron_b200.h
# ------------------------------------------------------------------------------------------------
def compact_labels(labels_a, classes_a, labels_b, classes_b, capacity, ignore_index=-1):
    """Labelled-row compaction on the device: (idx_a, lab_a, idx_b, lab_b, counts int32[8]); no host sync.
    counts[0:2] / counts[4:6] = (padded count, count) of a / b -- the row limits of the launches that follow."""
    L = init()
    n = labels_a.numel()
    dev = labels_a.device
    idx = torch.empty((2, 2, capacity), device=dev, dtype=torch.int64)
    counts = torch.empty(8, device=dev, dtype=torch.int32)
    check(L.vt_compact_labels(ptr(labels_a), classes_a, ptr(labels_b), classes_b, n, ignore_index, capacity, ptr(idx[0, 0]), ptr(idx[0, 1]),
                              ptr(idx[1, 0]), ptr(idx[1, 1]), ptr(counts), stream_ptr()), "vt_compact_labels")
    return idx[0, 0], idx[0, 1], idx[1, 0], idx[1, 1], counts


class row_limit:
    """`with row_limit(counts, which):` -- the launches inside clip their row extent on the device to counts[4 which] (padded) /
    counts[4 which + 1] (true count).  Thread-local in the library: forward and backward each bracket their own launches."""

    def __init__(self, counts, which=0):
        self.p = counts.data_ptr() + 16 * which
        self.ref = (counts, which)

    def __enter__(self):
        init().vt_set_row_limit(self.p)
        lib.ACTIVE_LIMIT.v = self.ref
        return self

    def __exit__(self, *exc):
        init().vt_set_row_limit(None)
        lib.ACTIVE_LIMIT.v = None
        return False


def seed_advance(state):
    """state: int64[2] on the device = (counter, seed); one tiny launch, the first node of a captured step."""
    check(init().vt_seed_advance(ptr(state), stream_ptr()), "vt_seed_advance")


def set_seed_source(state):
    """Dropout kernels XOR the device word state[1] into their by-value seed from now on (None: by-value seeds only)."""
    init().vt_set_seed_source(None if state is None else state.data_ptr() + 8)


def softmax_ce_fwd(logits, ncols, labels, double_softmax=False):
    L = init()
    rows = logits.size(0)
    dev = logits.device
    row_loss = torch.empty(rows, device=dev, dtype=F32)
    argmax = torch.empty(rows, device=dev, dtype=torch.int64)
    stats = torch.empty((rows, 2), device=dev, dtype=F32)
    check(L.vt_softmax_ce_fwd(ptr(logits), logits.stride(0), rows, ncols, ptr(labels), int(double_softmax), ptr(row_loss), ptr(argmax),
                              ptr(stats), stream_ptr()), "vt_softmax_ce_fwd")
    return row_loss, argmax, stats


def softmax_ce_bwd(logits, ncols, labels, stats, scale, scale_dev, double_softmax=False, ld_out=None):
    L = init()
    rows = logits.size(0)
    ld_out = roundup(ncols, 8) if ld_out is None else ld_out
    d = torch.empty((rows, ld_out), device=logits.device, dtype=BF16)
    check(L.vt_softmax_ce_bwd(ptr(logits), logits.stride(0), rows, ncols, ptr(labels), int(double_softmax), ptr(stats), scale,
                              ptr(scale_dev), ptr(d), ld_out, stream_ptr()), "vt_softmax_ce_bwd")
    return d


def action_head_fwd(pooled, W, bias, labels):
    L = init()
    B, Hd = pooled.shape
    A = W.size(0)
    scores = torch.empty((B, A), device=pooled.device, dtype=F32)
    row_loss = torch.empty(B, device=pooled.device, dtype=F32)
    argmax = torch.empty(B, device=pooled.device, dtype=torch.int64)
    check(L.vt_action_head_fwd(ptr(pooled), ptr(W), ptr(bias), B, Hd, A, ptr(labels), ptr(scores), ptr(row_loss), ptr(argmax), stream_ptr()),
          "vt_action_head_fwd")
    return scores, row_loss, argmax


def action_head_bwd(pooled, W, labels, scores, n_valid_dev, g_dev, dW, dbias):
    L = init()
    B, Hd = pooled.shape
    dpooled = torch.empty_like(pooled)
    check(L.vt_action_head_bwd(ptr(pooled), ptr(W), B, Hd, W.size(0), ptr(labels), ptr(scores), ptr(n_valid_dev), ptr(g_dev), ptr(dpooled),
                               ptr(dW), ptr(dbias), stream_ptr()), "vt_action_head_bwd")
    return dpooled


def pretrain_finalize(mlm, tok, act):
    """each of mlm/tok/act = (row_loss, argmax, labels); returns fp32 [8]."""
    L = init()
    out = torch.empty(8, device=act[0].device, dtype=F32)
    args = []
    for rl, am, lab in (mlm, tok, act):
        args += [ptr(rl), ptr(am), ptr(lab), rl.numel()]
    check(L.vt_pretrain_finalize(*args, ptr(out), stream_ptr()), "vt_pretrain_finalize")
    return out


def pretrain_finalize_counted(mlm, tok, act, counts):
    """pretrain_finalize with the MLM / token row counts read from compact_labels' counts on the device."""
    L = init()
    out = torch.empty(8, device=counts.device, dtype=F32)
    args = []
    for rl, am, lab in (mlm, tok):
        args += [ptr(rl), ptr(am), ptr(lab)]
    rl, am, lab = act
    args += [ptr(rl), ptr(am), ptr(lab), rl.numel()]
    check(L.vt_pretrain_finalize_counted(*args, ptr(counts), ptr(out), stream_ptr()), "vt_pretrain_finalize_counted")
    return out


def tanh_bwd(dy, y):
    L = init()
    dx = torch.empty(y.shape, device=y.device, dtype=BF16)
    check(L.vt_tanh_bwd(ptr(dy), ptr(y), y.numel(), ptr(dx), stream_ptr()), "vt_tanh_bwd")
    return dx


def to_f32(x):
    L = init()
    y = torch.empty(x.shape, device=x.device, dtype=F32)
    check(L.vt_cast_bf16_f32(ptr(x), x.numel(), ptr(y), stream_ptr()), "vt_cast_bf16_f32")
    return y


def gelu_bwd(dy, pre):
    L = init()
    dx = torch.empty_like(pre)
    check(L.vt_gelu_bwd(ptr(dy), ptr(pre), pre.numel(), ptr(dx), stream_ptr()), "vt_gelu_bwd")
    return dx


def to_bf16(x, addend=None, out=None):
    L = init()
    y = torch.empty(x.shape, device=x.device, dtype=BF16) if out is None else out
    check(L.vt_cast_f32_bf16_add(ptr(x), ptr(addend), x.numel(), ptr(y), stream_ptr()), "vt_cast_f32_bf16_add")
    return y
