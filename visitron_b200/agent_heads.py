"""The fine-tuning heads of the reference's agents on the B200 kernels.

    SoftDotAttention          vs/agent_models.py:313-357 (query_dim, ctx_dim) == tb/agent_models.py:238-274 (dim)
    candidate scorer          AttnDecoderLSTM.candidate_att_layer, vs/agent_models.py:382,426 + vs/agent.py:396-408
    turn-based action head    attention_layer + decoder2action, tb/agent_models.py:297-318 + tb/agent.py:150-152,316-327
    question-asking head      question_linear, vs/agent_models.py:455-459,503 + vs/classifier/agent.py:158-163,493-503

`SoftDotAttention` keeps the reference's constructor, parameter names (`linear_in.weight`, `linear_out.weight`) and
`forward(h, context, mask=None, output_tilde=True, output_prob=True)`, so it can replace the class in agent_models.py;
`masked_cross_entropy`, `bce_with_logits_masked_mean` replace the loss lines of the agents; `candidate_scores_ce`,
`turn_based_head` and `qa_head` are the fused compositions, with the argument order of the oracle's functions of the
same names.  Everything is fp32 (the reference's dtype) so that the action argmax agrees exactly; the arithmetic runs in
libvisitron_b200.so (csrc/agent_heads.cu) -- there is no eager fallback.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import lib
from .lib import check, init, ptr, stream_ptr

F32 = torch.float32


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != F32:
        raise lib.VtError("the fine-tuning heads run in fp32 (the reference's dtype); got %s" % t.dtype)
    return t.contiguous()


def _u8(mask: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if mask is None:
        return None
    if mask.dtype == torch.bool:
        return mask.contiguous().view(torch.uint8)
    if mask.dtype == torch.uint8:
        return mask.contiguous()
    return (mask != 0).contiguous().view(torch.uint8)


# ------------------------------------------------------------------------------------------------
# y = act([x1 ; x2] W^T + b)
# ------------------------------------------------------------------------------------------------
class _SmallLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x1, x2, W, bias, act: int):
        lib.require_device(x1)
        L = init()
        x1 = _f32c(x1)
        x2 = None if x2 is None else _f32c(x2)
        W = _f32c(W)
        bias = None if bias is None else _f32c(bias)
        B, K1 = x1.shape
        K2 = 0 if x2 is None else x2.size(1)
        N = W.size(0)
        if W.size(1) != K1 + K2:
            raise lib.VtError("small_linear: weight is [%d,%d] but the input has %d + %d columns" % (N, W.size(1), K1, K2))
        y = torch.empty((B, N), device=x1.device, dtype=F32)
        check(L.vt_small_linear_fwd(ptr(x1), x1.stride(0), K1, ptr(x2), 0 if x2 is None else x2.stride(0), K2, ptr(W), W.stride(0),
                                    ptr(bias), act, ptr(y), y.stride(0), B, N, stream_ptr()), "vt_small_linear_fwd")
        ctx.save_for_backward(x1, x2, W, y)
        ctx.act, ctx.has_bias = act, bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        L = init()
        x1, x2, W, y = ctx.saved_tensors
        dy = _f32c(dy)
        B, K1 = x1.shape
        K2 = 0 if x2 is None else x2.size(1)
        N = W.size(0)
        need = ctx.needs_input_grad
        dx1 = torch.empty_like(x1) if need[0] else None
        dx2 = torch.empty_like(x2) if (x2 is not None and need[1]) else None
        dW = torch.empty_like(W) if need[2] else None
        db = torch.empty(N, device=W.device, dtype=F32) if (ctx.has_bias and need[3]) else None
        if db is not None and dW is None:
            dW = torch.empty_like(W)  # the bias gradient is produced by the weight-gradient kernel
        nws = int(L.vt_small_linear_bwd_workspace(B, N, K1, K2)) if (dx1 is not None or dx2 is not None) else 0
        ws = torch.empty(nws, device=W.device, dtype=F32) if nws else None
        check(L.vt_small_linear_bwd(ptr(dy), dy.stride(0), ptr(y), y.stride(0), ctx.act, ptr(x1), x1.stride(0), K1, ptr(x2),
                                    0 if x2 is None else x2.stride(0), K2, ptr(W), W.stride(0), ptr(dx1), K1, ptr(dx2), K2, ptr(dW),
                                    W.size(1), ptr(db), B, N, ptr(ws), stream_ptr()), "vt_small_linear_bwd")
        return dx1, dx2, (dW if need[2] else None), db, None


def small_linear(x, W, bias=None, act: str = "none", x2=None):
    """act([x ; x2] W^T + bias) for a decoder step's batch; act in {"none", "tanh"}."""
    return _SmallLinearFn.apply(x, x2, W, bias, {"none": 0, "tanh": 1}[act])


# ------------------------------------------------------------------------------------------------
# SoftDotAttention core
# ------------------------------------------------------------------------------------------------
class _SoftDotFn(torch.autograd.Function):
    """(context, target, mask) -> (weighted_context, attn, logit); the returned logit is masked like the reference's."""

    @staticmethod
    def forward(ctx, context, target, mask):
        lib.require_device(context)
        L = init()
        context, target = _f32c(context), _f32c(target)
        B, Ln, D = context.shape
        if tuple(target.shape) != (B, D):
            raise lib.VtError("SoftDotAttention: linear_in(h) is %s but the context is %s" % (tuple(target.shape), tuple(context.shape)))
        m8 = _u8(mask)
        if m8 is not None and tuple(m8.shape) != (B, Ln):
            raise lib.VtError("SoftDotAttention: mask must be [batch, seq_len]")
        dev = context.device
        logit = torch.empty((B, Ln), device=dev, dtype=F32)
        attn = torch.empty((B, Ln), device=dev, dtype=F32)
        weighted = torch.empty((B, D), device=dev, dtype=F32)
        P = int(L.vt_softdot_splits(B, Ln))
        ws = torch.empty(B * P * (2 + D), device=dev, dtype=F32)
        check(L.vt_softdot_fwd(ptr(context), ptr(target), ptr(m8), B, Ln, D, ptr(logit), ptr(attn), ptr(weighted), ptr(ws), stream_ptr()),
              "vt_softdot_fwd")
        ctx.save_for_backward(context, target, attn)
        ctx.m8 = m8
        ctx.set_materialize_grads(False)
        return weighted, attn, logit

    @staticmethod
    def backward(ctx, dweighted, dattn, dlogit):
        L = init()
        context, target, attn = ctx.saved_tensors
        B, Ln, D = context.shape
        dev = context.device
        if dweighted is None and dattn is None and dlogit is None:
            return None, None, None
        dweighted = None if dweighted is None else _f32c(dweighted)
        dattn = None if dattn is None else _f32c(dattn)
        dlogit = None if dlogit is None else _f32c(dlogit)
        dctx = torch.empty_like(context) if ctx.needs_input_grad[0] else None
        dtarget = torch.empty_like(target)
        P = int(L.vt_softdot_splits(B, Ln))
        ws = torch.empty(B * Ln + B * P * D, device=dev, dtype=F32)
        check(L.vt_softdot_bwd(ptr(context), ptr(target), ptr(ctx.m8), ptr(attn), ptr(dweighted), ptr(dattn), ptr(dlogit), B, Ln, D,
                               ptr(dctx), ptr(dtarget), ptr(ws), stream_ptr()), "vt_softdot_bwd")
        return dctx, dtarget, None


class SoftDotAttention(nn.Module):
    """Drop-in for vs/agent_models.py:313-357 (`SoftDotAttention(query_dim, ctx_dim)`) and tb/agent_models.py:238-274
    (`SoftDotAttention(dim)`): same parameters, same forward signature and return values."""

    def __init__(self, query_dim, ctx_dim=None):
        super().__init__()
        ctx_dim = query_dim if ctx_dim is None else ctx_dim
        self.linear_in = nn.Linear(query_dim, ctx_dim, bias=False)
        self.sm = nn.Softmax(dim=1)
        self.linear_out = nn.Linear(query_dim + ctx_dim, query_dim, bias=False)
        self.tanh = nn.Tanh()

    def forward(self, h, context, mask=None, output_tilde=True, output_prob=True):
        target = small_linear(h, self.linear_in.weight)
        weighted, attn, logit = _SoftDotFn.apply(context, target, mask)
        second = attn if output_prob else logit
        if output_tilde:
            h_tilde = small_linear(weighted, self.linear_out.weight, act="tanh", x2=h)
            return h_tilde, second
        return weighted, second


# ------------------------------------------------------------------------------------------------
# losses
# ------------------------------------------------------------------------------------------------
class _HeadCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, valid_len, forbid0, target, ignore_index: int, forbid_col: int = 4, empty_is_zero: bool = False):
        lib.require_device(logits)
        L = init()
        logits = _f32c(logits)
        B, C = logits.shape
        dev = logits.device
        vl = None if valid_len is None else valid_len.to(device=dev, dtype=torch.int64).contiguous()
        f0 = _u8(forbid0)
        target = target.to(device=dev, dtype=torch.int64).contiguous()
        masked = torch.empty_like(logits)
        lse = torch.empty(B, device=dev, dtype=F32)
        argmax = torch.empty(B, device=dev, dtype=torch.int64)
        out3 = torch.empty(3, device=dev, dtype=F32)
        check(L.vt_head_ce_fwd(ptr(logits), logits.stride(0), B, C, ptr(vl), ptr(f0), forbid_col, int(empty_is_zero), ptr(target), ignore_index, ptr(masked), C, ptr(lse),
                               ptr(argmax), ptr(out3), stream_ptr()), "vt_head_ce_fwd")
        ctx.save_for_backward(masked, target, lse, out3)
        ctx.ignore_index = ignore_index
        ctx.mark_non_differentiable(masked, argmax)
        return out3[0], masked, argmax

    @staticmethod
    def backward(ctx, g, _gm, _ga):
        L = init()
        masked, target, lse, out3 = ctx.saved_tensors
        B, C = masked.shape
        g = g.reshape(1).to(F32).contiguous()
        d = torch.empty_like(masked)
        check(L.vt_head_ce_bwd(ptr(masked), C, B, C, ptr(target), ctx.ignore_index, ptr(lse), ptr(out3), ptr(g), ptr(d), C, stream_ptr()),
              "vt_head_ce_bwd")
        return d, None, None, None, None, None, None


FORWARD_ACTION = 4   # tb/agent.py:71-80: model_actions = [left, right, up, down, forward, <end>, <start>, <ignore>]


def masked_cross_entropy(logit, target, valid_len=None, forbid=None, ignore_index: int = -100, forbid_col: int = FORWARD_ACTION,
                         empty_is_zero: bool = False):
    """masked_fill(-inf) of the impossible actions + CrossEntropyLoss(ignore_index) + argmax in one pass.

    valid_len [B]: columns >= valid_len[b] are impossible (candidate_leng, vs/agent.py:396-403);
    forbid [B] bool: column `forbid_col` (model_actions.index("forward") = 4) is impossible (tb/agent.py:316-318).
    Returns (loss, masked_logit, argmax)."""
    return _HeadCEFn.apply(logit, valid_len, forbid, target, int(ignore_index), int(forbid_col), bool(empty_is_zero))


class _BCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, target, valid, pos_weight: float):
        lib.require_device(q)
        L = init()
        q = _f32c(q.reshape(-1))
        B = q.numel()
        target = target.to(device=q.device, dtype=F32).reshape(-1).contiguous()
        v8 = _u8(valid)
        per = torch.empty(B, device=q.device, dtype=F32)
        out2 = torch.empty(2, device=q.device, dtype=F32)
        check(L.vt_bce_logits_fwd(ptr(q), ptr(target), ptr(v8), B, float(pos_weight), ptr(per), ptr(out2), stream_ptr()), "vt_bce_logits_fwd")
        ctx.save_for_backward(q, target, out2)
        ctx.v8, ctx.pw = v8, float(pos_weight)
        ctx.mark_non_differentiable(per)
        return out2[0], per

    @staticmethod
    def backward(ctx, g, _gp):
        L = init()
        q, target, out2 = ctx.saved_tensors
        g = g.reshape(1).to(F32).contiguous()
        dq = torch.empty_like(q)
        check(L.vt_bce_logits_bwd(ptr(q), ptr(target), ptr(ctx.v8), q.numel(), ctx.pw, ptr(out2), ptr(g), ptr(dq), stream_ptr()),
              "vt_bce_logits_bwd")
        return dq, None, None, None


def bce_with_logits_masked_mean(q, target, valid=None, pos_weight: float = 5.0):
    """BCEWithLogitsLoss(pos_weight, reduction="none") * valid, sum / count (0 when nothing is valid).  Returns (loss, per_row)."""
    shape = q.shape
    loss, per = _BCEFn.apply(q.reshape(-1), target, valid, float(pos_weight))
    return loss, per.view(shape)


# ------------------------------------------------------------------------------------------------
# fused compositions, argument order of the oracle's functions of the same names
# ------------------------------------------------------------------------------------------------
def candidate_scores_ce(w_in, h_tilde, cand_feat, cand_len, target, ignore_index: int = -100):
    """vs/agent_models.py:426 + vs/agent.py:396-408 -> (masked logit [B,C], loss, argmax)."""
    t = small_linear(h_tilde, w_in)
    _, _, logit = _SoftDotFn.apply(cand_feat, t, None)
    loss, masked, am = masked_cross_entropy(logit, target, valid_len=cand_len, ignore_index=ignore_index)
    return masked, loss, am


def turn_based_head(w_in, w_out, w_a, b_a, h, ctx, ctx_mask, no_forward, target, ignore_index: int = 7):
    """tb/agent_models.py:253-274,297,316-318 + tb/agent.py:150-152,316-327 -> (h_tilde, masked logit [B,6], loss, argmax)."""
    t = small_linear(h, w_in)
    weighted, _, _ = _SoftDotFn.apply(ctx, t, ctx_mask)
    h_tilde = small_linear(weighted, w_out, act="tanh", x2=h)
    logit = small_linear(h_tilde, w_a, b_a)
    loss, masked, am = masked_cross_entropy(logit, target, forbid=no_forward, ignore_index=ignore_index)
    return h_tilde, masked, loss, am


def qa_head(w1, b1, w2, b2, h_tilde, target, valid, pos_weight: float = 5.0):
    """vs/agent_models.py:455-459,503 + vs/classifier/agent.py:158-163,493-503 -> (logit [B,1], loss)."""
    z = small_linear(h_tilde, w1, b1, act="tanh")
    q = small_linear(z, w2, b2)
    loss, _ = bce_with_logits_masked_mean(q, target, valid, pos_weight)
    return q, loss
