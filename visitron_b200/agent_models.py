"""The recurrent middle of the reference's agents on the B200 kernels: same class names, constructors,
parameter names and forward() signatures as tasks/viewpoint_select/agent_models.py and tasks/turn_based/agent_models.py.

    OscarEncoder                        vs/agent_models.py:192-310 == tb/agent_models.py:117-235
    AttnDecoderLSTM (viewpoint select)  vs/agent_models.py:360-428   -> AttnDecoderLSTM
    AttnDecoderLSTM (turn based)        tb/agent_models.py:277-319   -> TurnAttnDecoderLSTM
    SoftDotAttention                    visitron_b200/agent_heads.py

The sub-modules (nn.LSTM, nn.LSTMCell, nn.Linear, nn.Embedding) are PARAMETER CONTAINERS with the reference's state_dict
keys (``lstm.weight_ih_l0`` ...); they are never called.  The arithmetic runs in libvisitron_b200.so: the LSTM's input
projection over all 511 positions is one tcgen05 GEMM on the encoder's bf16 output (no fp32 round trip), the recurrence is
the cluster kernel of csrc/lstm.cu, the decoder step is a chain of the step-sized kernels of csrc/agent_heads.cu plus the
LSTMCell / dropout / embedding kernels of csrc/lstm.cu.  fp32 state and logits (the reference's dtype), so that the action
argmax agrees.  There is no eager fallback.
"""
from __future__ import annotations

import warnings
from typing import Optional

import torch
from torch import nn

from . import lib, modeling, ops
from .agent_heads import SoftDotAttention, small_linear
from .lib import check, init, ptr, stream_ptr

F32, BF16 = torch.float32, torch.bfloat16
SITE_AGENT = 2000     # dropout-site ids of the agent modules (the encoder uses 4 l + k and 1000 / 1001)


# ------------------------------------------------------------------------------------------------
# small autograd nodes over the C ABI
# ------------------------------------------------------------------------------------------------
class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p: float, seed: int, site: int):
        x = x.contiguous()
        y = torch.empty_like(x)
        check(init().vt_dropout_f32(ptr(x), x.numel(), p, seed, site, ptr(y), stream_ptr()), "vt_dropout_f32")
        ctx.p, ctx.seed, ctx.site = p, seed, site
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        check(init().vt_dropout_f32(ptr(dy), dy.numel(), ctx.p, ctx.seed, ctx.site, ptr(dx), stream_ptr()), "vt_dropout_f32")
        return dx, None, None, None


class _Drop:
    """nn.Dropout of an agent module: identity in eval mode; in training one seed per forward call of the owning module and a
    site id per call site, so that forward and backward regenerate the same mask."""

    def __init__(self, p: float):
        self.p = float(p)
        self.seed = 0
        self.n = 0

    def begin(self, training: bool, site0: int):
        self.active = training and self.p > 0
        self.seed = modeling._draw_seed() if self.active else 0
        self.n = site0

    def __call__(self, x):
        if not self.active:
            return x
        if x.dtype != F32:
            raise lib.VtError("agent dropout runs in fp32 (the reference's dtype); got %s" % x.dtype)
        self.n += 1
        return _DropoutFn.apply(x, self.p, self.seed, self.n)


class _LSTMCellFn(torch.autograd.Function):
    """(ga, gb, c0) -> (h1, c1), ga / gb = the two halves of the gate pre-activations (biases included)."""

    @staticmethod
    def forward(ctx, ga, gb, c0):
        lib.require_device(ga)
        ga, gb, c0 = ga.contiguous(), gb.contiguous(), c0.contiguous()
        B, Hh = c0.shape
        act = torch.empty_like(ga)
        h1, c1 = torch.empty_like(c0), torch.empty_like(c0)
        check(init().vt_lstm_cell_fwd(ptr(ga), ptr(gb), ptr(c0), B, Hh, ptr(act), ptr(h1), ptr(c1), stream_ptr()), "vt_lstm_cell_fwd")
        ctx.save_for_backward(act, c0, c1)
        ctx.set_materialize_grads(False)
        return h1, c1

    @staticmethod
    def backward(ctx, dh1, dc1):
        act, c0, c1 = ctx.saved_tensors
        B, Hh = c0.shape
        if dh1 is None and dc1 is None:
            return None, None, None
        dh1 = None if dh1 is None else dh1.contiguous()
        dc1 = None if dc1 is None else dc1.contiguous()
        dg = torch.empty_like(act)
        dc0 = torch.empty_like(c0)
        check(init().vt_lstm_cell_bwd(ptr(act), ptr(c0), ptr(c1), ptr(dh1), ptr(dc1), B, Hh, ptr(dg), ptr(dc0), stream_ptr()), "vt_lstm_cell_bwd")
        return dg, dg, dc0


class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, W, idx):
        lib.require_device(W)
        idx = idx.reshape(-1).contiguous().long()
        V, D = W.shape
        y = torch.empty((idx.numel(), D), device=W.device, dtype=F32)
        check(init().vt_embedding_f32_fwd(ptr(W.contiguous()), ptr(idx), idx.numel(), V, D, ptr(y), stream_ptr()), "vt_embedding_f32_fwd")
        ctx.idx, ctx.shape = idx, (V, D)
        return y

    @staticmethod
    def backward(ctx, dy):
        V, D = ctx.shape
        dy = dy.contiguous()
        dW = torch.empty((V, D), device=dy.device, dtype=F32)
        check(init().vt_embedding_f32_bwd(ptr(dy), ptr(ctx.idx), ctx.idx.numel(), V, D, ptr(dW), stream_ptr()), "vt_embedding_f32_bwd")
        return dW, None


class _LSTMSeqFn(torch.autograd.Function):
    """nn.LSTM(768 -> 512) over the encoder's bf16 output rows [B*T, 768] (row b*T + t): (ctx [B,T,512], h_last, c_last)."""

    @staticmethod
    def forward(ctx, x, lengths, B: int, T: int, w_ih, w_hh, b_ih, b_hh):
        L = init()
        dev = x.device
        G, Hh = w_hh.shape
        wih_b = torch.empty(tuple(w_ih.shape), device=dev, dtype=BF16)
        ops.cast_weight(w_ih.detach(), wih_b)
        bias = (b_ih.detach() + b_hh.detach()).contiguous()
        gates = lib.gemm(x, wih_b, bias=bias, out_dtype=F32)              # [B*T, 2048]: every step's input projection at once
        out = torch.empty((B, T, Hh), device=dev, dtype=F32)
        c_all = torch.empty((B, T, Hh), device=dev, dtype=F32)
        hprev = torch.empty((B * T, Hh), device=dev, dtype=BF16)
        h_last = torch.empty((B, Hh), device=dev, dtype=F32)
        c_last = torch.empty((B, Hh), device=dev, dtype=F32)
        w_hh_c = w_hh.detach().contiguous()
        check(L.vt_lstm_seq_fwd(ptr(gates), ptr(w_hh_c), ptr(lengths), B, T, Hh, ptr(out), ptr(c_all), ptr(hprev), ptr(h_last), ptr(c_last),
                                stream_ptr()), "vt_lstm_seq_fwd")
        ctx.saved = (x, lengths, wih_b, w_hh_c, gates, c_all, hprev)
        ctx.dims = (B, T, G, Hh)
        ctx.set_materialize_grads(False)
        return out, h_last, c_last

    @staticmethod
    def backward(ctx, dctx, dh_last, dc_last):
        L = init()
        x, lengths, wih_b, w_hh, gates, c_all, hprev = ctx.saved
        ctx.saved = None
        B, T, G, Hh = ctx.dims
        dev = x.device
        dctx = torch.zeros((B, T, Hh), device=dev, dtype=F32) if dctx is None else dctx.contiguous().float()
        dh_last = None if dh_last is None else dh_last.contiguous()
        dc_last = None if dc_last is None else dc_last.contiguous()
        dA = torch.empty((B * T, G), device=dev, dtype=BF16)
        check(L.vt_lstm_seq_bwd(ptr(gates), ptr(c_all), ptr(w_hh), ptr(lengths), ptr(dctx), ptr(dh_last), ptr(dc_last), B, T, Hh, ptr(dA),
                                stream_ptr()), "vt_lstm_seq_bwd")
        need = ctx.needs_input_grad
        dw_ih = dw_hh = db = dx = None
        if need[4]:
            dw_ih = torch.zeros((G, x.size(1)), device=dev, dtype=F32)
            lib.gemm(dA, x, a_major=1, b_major=1, out=dw_ih, accumulate=True)
        if need[5]:
            dw_hh = torch.zeros((G, Hh), device=dev, dtype=F32)
            lib.gemm(dA, hprev, a_major=1, b_major=1, out=dw_hh, accumulate=True)
        if need[6] or need[7]:
            db = ops.colsum(dA, torch.zeros(G, device=dev, dtype=F32), accumulate=False)
        if need[0]:
            dx = lib.gemm(dA, wih_b, b_major=1)                           # bf16 [B*T, 768]: straight into the encoder's backward
        return dx, None, None, None, dw_ih, dw_hh, db, db


def lstm_cell(x1, x2, h, c, w_ih, w_hh, b_ih, b_hh):
    """nn.LSTMCell on the concatenation [x1 ; x2] (x2 may be None) -> (h1, c1)."""
    ga = small_linear(x1, w_ih, b_ih, x2=x2)
    gb = small_linear(h, w_hh, b_hh)
    return _LSTMCellFn.apply(ga, gb, c)


# ------------------------------------------------------------------------------------------------
# OscarEncoder
# ------------------------------------------------------------------------------------------------
class OscarEncoder(nn.Module):
    """vs/agent_models.py:192-310 == tb/agent_models.py:117-235: the Oscar encoder (text only), a 1-layer LSTM over all its
    positions, and the decoder's initial state."""

    def __init__(self, args, bert, hidden_size, decoder_hidden_size, dropout_ratio, bidirectional=False, num_layers=1, reverse_input=False):
        super().__init__()
        if bidirectional or num_layers != 1:
            raise NotImplementedError("the cluster LSTM kernel is the 1-layer unidirectional nn.LSTM every run script uses")
        if reverse_input:
            raise NotImplementedError("reverse_input is not supported (no run script sets it)")
        if hidden_size != 512:
            raise NotImplementedError("hidden_size %d outside the supported envelope (512 = --rnn_dim of every run script)" % hidden_size)
        self.transformer_hidden_size = 768
        self.reverse_input = reverse_input
        self.dec_hidden_size = decoder_hidden_size
        self.args = args
        self.bert = bert
        self.hidden_size = hidden_size
        self.drop = nn.Dropout(p=dropout_ratio)
        self.num_directions = 1
        self.num_layers = num_layers
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")      # "dropout option adds dropout after all but last recurrent layer": the reference's call
            self.lstm = nn.LSTM(self.transformer_hidden_size, self.hidden_size, self.num_layers, batch_first=True, dropout=dropout_ratio,
                                bidirectional=False)
        self.encoder_lstm2decoder_ht = nn.Linear(hidden_size, decoder_hidden_size)
        self.encoder_lstm2decoder_ct = nn.Linear(hidden_size, decoder_hidden_size)
        self._drop = _Drop(dropout_ratio)

    def _vt_set_reducer(self, reducer):      # parallel.distribute() reaches the encoder through self.bert
        pass

    def forward(self, inputs, lengths, mask, position_ids=None, token_type_ids=None):
        if not hasattr(self.bert, "_vt_forward"):
            raise lib.VtError("OscarEncoder needs the visitron_b200 BertImgModelwithLocationEmbeds as `bert`")
        att_mask = ~mask                                                 # vs/agent_models.py:268 (uint8: 255)
        seq, _, step = self.bert._vt_forward(inputs, token_type_ids=token_type_ids, attention_mask=att_mask, position_ids=position_ids)
        B, T = step.B, step.S
        if torch.is_tensor(lengths):
            max_len = int(lengths.max()) if not lengths.is_cuda else T   # a device tensor of lengths is not synced on
            len_dev = lengths.to(device=seq.device, dtype=torch.int64, non_blocking=True)
        else:
            max_len = int(max(lengths))
            len_dev = torch.tensor(list(lengths), dtype=torch.int64).to(seq.device, non_blocking=True)
        l = self.lstm
        ctx, h_t, c_t = _LSTMSeqFn.apply(seq, len_dev, B, T, l.weight_ih_l0, l.weight_hh_l0, l.bias_ih_l0, l.bias_hh_l0)
        decoder_init = small_linear(h_t, self.encoder_lstm2decoder_ht.weight, self.encoder_lstm2decoder_ht.bias, act="tanh")
        if self.hidden_size * self.num_directions != self.dec_hidden_size:
            c_t = small_linear(c_t, self.encoder_lstm2decoder_ct.weight, self.encoder_lstm2decoder_ct.bias)
        if max_len < T:
            ctx = ctx[:, :max_len]                                       # pad_packed_sequence pads to the longest sample
        self._drop.begin(self.training, SITE_AGENT)
        ctx = self._drop(ctx)
        return ctx, decoder_init, c_t


# ------------------------------------------------------------------------------------------------
# decoders
# ------------------------------------------------------------------------------------------------
class AttnDecoderLSTM(nn.Module):
    """vs/agent_models.py:360-428: one step of the viewpoint-selection decoder."""

    def __init__(self, angle_feat_size, embedding_size, hidden_size, dropout_ratio, feature_size=2048 + 4):
        super().__init__()
        self.embedding_size = embedding_size
        self.feature_size = feature_size
        self.hidden_size = hidden_size
        self.embedding = nn.Sequential(nn.Linear(angle_feat_size, self.embedding_size), nn.Tanh())
        self.drop = nn.Dropout(p=dropout_ratio)
        self.lstm = nn.LSTMCell(embedding_size + feature_size, hidden_size)
        self.feat_att_layer = SoftDotAttention(hidden_size, feature_size)
        self.attention_layer = SoftDotAttention(hidden_size, hidden_size)
        self.candidate_att_layer = SoftDotAttention(hidden_size, feature_size)
        self._drop = _Drop(dropout_ratio)
        self._calls = 0

    def forward(self, action, feature, cand_feat, h_0, prev_h1, c_0, ctx, ctx_mask=None):
        self._calls += 1
        drop = self._drop
        drop.begin(self.training, SITE_AGENT + 16 * (self._calls & 0xffff))
        action_embeds = drop(small_linear(action, self.embedding[0].weight, self.embedding[0].bias, act="tanh"))
        prev_h1_drop = drop(prev_h1)
        attn_feat, _ = self.feat_att_layer(prev_h1_drop, feature, output_tilde=False)
        c = self.lstm
        h_1, c_1 = lstm_cell(action_embeds, attn_feat, prev_h1, c_0, c.weight_ih, c.weight_hh, c.bias_ih, c.bias_hh)
        h_1_drop = drop(h_1)
        h_tilde, _alpha = self.attention_layer(h_1_drop, ctx, ctx_mask)
        h_tilde_drop = drop(h_tilde)
        _, logit = self.candidate_att_layer(h_tilde_drop, cand_feat, output_prob=False)
        return h_1, c_1, logit, h_tilde


class TurnAttnDecoderLSTM(nn.Module):
    """tb/agent_models.py:277-319 (there also called AttnDecoderLSTM): one step of the turn-based decoder."""

    def __init__(self, input_action_size, output_action_size, embedding_size, hidden_size, dropout_ratio, feature_size=2048):
        super().__init__()
        self.embedding_size = embedding_size
        self.feature_size = feature_size
        self.hidden_size = hidden_size
        self.embedding = nn.Embedding(input_action_size, embedding_size)
        self.drop = nn.Dropout(p=dropout_ratio)
        self.lstm = nn.LSTMCell(embedding_size + feature_size, hidden_size)
        self.attention_layer = SoftDotAttention(hidden_size)
        self.decoder2action = nn.Linear(hidden_size, output_action_size)
        self._drop = _Drop(dropout_ratio)
        self._calls = 0

    def forward(self, action, feature, h_0, c_0, ctx, ctx_mask=None):
        self._calls += 1
        drop = self._drop
        drop.begin(self.training, SITE_AGENT + 16 * (self._calls & 0xffff))
        action_embeds = _EmbeddingFn.apply(self.embedding.weight, action)             # [b, 1] -> [b, embedding_size]
        concat_input = torch.cat((action_embeds, feature.float()), 1)                 # data movement only
        dropped = drop(concat_input)
        c = self.lstm
        h_1, c_1 = lstm_cell(dropped, None, h_0, c_0, c.weight_ih, c.weight_hh, c.bias_ih, c.bias_hh)
        h_1_drop = drop(h_1)
        h_tilde, alpha = self.attention_layer(h_1_drop, ctx, ctx_mask)
        logit = small_linear(h_tilde, self.decoder2action.weight, self.decoder2action.bias)
        return h_1, c_1, alpha, logit
