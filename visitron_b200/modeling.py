"""The reference's model classes on the B200 kernels: same class names, constructor ``(config)``,
``forward()`` signatures, return tuples and ``state_dict`` keys, so they drop
into ``MODEL_CLASS`` of tasks/viewpoint_select/model_utils.py:15-26 and into the training loops.

    BertImgModelwithLocationEmbeds   vs/encoder.py:161-303   (== tb/encoder.py:159-301)
    PreTrainOscar                    vs/encoder.py:306-441
    CaptionBert{Encoder,Layer,Attention,SelfAttention}        oscar/modeling_bert.py:26-169
    BertEmbeddings, BertLayerNorm, BertSelfOutput, BertIntermediate, BertOutput, BertPooler,
    BertOnlyMLMHead (+ BertLMPredictionHead, BertPredictionHeadTransform)   hf/modeling_bert.py
    NextActionPrediction             vs/encoder.py:142-158

The sub-modules are parameter containers with the reference's names; the arithmetic runs in two
autograd nodes -- the whole encoder, and the pre-training heads -- whose forward/backward are
sequences of calls into libvisitron_b200.so (include/visitron_b200.h).  Activations are bf16,
accumulation / statistics / losses / master weights / gradients are fp32.  There is no eager
fallback: without the library or off a B200 every forward raises.
"""
from __future__ import annotations

import contextlib
import logging
from typing import List, Optional

import torch
from torch import nn

from . import lib, ops
from .flat import FlatStore
from .lib import ACT_GELU, ACT_GELU_BWD, ACT_GELU_SAVE_GRAD, ACT_MUL_AUX, ACT_TANH
from .pretrained import PreTrainedModel, check_envelope

logger = logging.getLogger(__name__)

BF16, F32 = torch.bfloat16, torch.float32
H = 768
LOC_DIM = 128                      # vs/encoder.py:179 hard-codes 128
STREAM_EMB, STREAM_IMG = 1000, 1001  # dropout-site ids; encoder layer l uses 4l (probs), 4l+1 (attn out), 4l+2 (ffn out)


# ------------------------------------------------------------------------------------------------
# parameter containers (names == the reference's state_dict keys)
# ------------------------------------------------------------------------------------------------
class _FusedOnly(nn.Module):
    """Sub-modules hold parameters; they run fused inside the encoder / head nodes."""

    def forward(self, *a, **k):
        raise NotImplementedError("%s runs fused inside its parent model on the CUDA path and cannot be called on its own"
                                  % type(self).__name__)


class BertLayerNorm(_FusedOnly):
    """hf/modeling_bert.py:227-240 (TF style: biased variance, eps inside the sqrt)."""

    def __init__(self, hidden_size, eps=1e-12):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.bias = nn.Parameter(torch.zeros(hidden_size))
        self.variance_epsilon = eps


class BertEmbeddings(_FusedOnly):
    """hf/modeling_bert.py:243-271."""

    def __init__(self, config):
        super().__init__()
        self.word_embeddings = nn.Embedding(config.vocab_size, config.hidden_size, padding_idx=0)
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_size)
        self.token_type_embeddings = nn.Embedding(config.type_vocab_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)


class CaptionBertSelfAttention(_FusedOnly):
    """oscar/modeling_bert.py:26-79 (base hf/modeling_bert.py:274-296)."""

    def __init__(self, config):
        super().__init__()
        if config.hidden_size % config.num_attention_heads != 0:
            raise ValueError("The hidden size (%d) is not a multiple of the number of attention heads (%d)"
                             % (config.hidden_size, config.num_attention_heads))
        self.output_attentions = getattr(config, "output_attentions", False)
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.all_head_size = self.num_attention_heads * self.attention_head_size
        self.query = nn.Linear(config.hidden_size, self.all_head_size)
        self.key = nn.Linear(config.hidden_size, self.all_head_size)
        self.value = nn.Linear(config.hidden_size, self.all_head_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)


class BertSelfOutput(_FusedOnly):
    """hf/modeling_bert.py:334-345."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)


class CaptionBertAttention(_FusedOnly):
    """oscar/modeling_bert.py:82-98."""

    def __init__(self, config):
        super().__init__()
        self.self = CaptionBertSelfAttention(config)
        self.output = BertSelfOutput(config)

    def prune_heads(self, heads):
        raise NotImplementedError("head pruning changes the 12 x 64 head geometry the CUDA kernels are specialised for")


class BertIntermediate(_FusedOnly):
    """hf/modeling_bert.py:378-390."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.intermediate_size)


class BertOutput(_FusedOnly):
    """hf/modeling_bert.py:393-404."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.intermediate_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)


class CaptionBertLayer(_FusedOnly):
    """oscar/modeling_bert.py:101-124."""

    def __init__(self, config):
        super().__init__()
        self.attention = CaptionBertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)


class CaptionBertEncoder(_FusedOnly):
    """oscar/modeling_bert.py:127-169."""

    def __init__(self, config):
        super().__init__()
        self.output_attentions = getattr(config, "output_attentions", False)
        self.output_hidden_states = getattr(config, "output_hidden_states", False)
        self.layer = nn.ModuleList([CaptionBertLayer(config) for _ in range(config.num_hidden_layers)])


class BertPooler(_FusedOnly):
    """hf/modeling_bert.py:455-467."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.activation = nn.Tanh()


class BertPredictionHeadTransform(_FusedOnly):
    """hf/modeling_bert.py:470-486."""

    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=config.layer_norm_eps)


class BertLMPredictionHead(_FusedOnly):
    """hf/modeling_bert.py:489-503: decoder weight tied to the word embeddings, separate output bias."""

    def __init__(self, config):
        super().__init__()
        self.transform = BertPredictionHeadTransform(config)
        self.decoder = nn.Linear(config.hidden_size, config.vocab_size, bias=False)
        self.bias = nn.Parameter(torch.zeros(config.vocab_size))


class BertOnlyMLMHead(_FusedOnly):
    """hf/modeling_bert.py:506-513."""

    def __init__(self, config):
        super().__init__()
        self.predictions = BertLMPredictionHead(config)


class NextActionPrediction(_FusedOnly):
    """vs/encoder.py:142-158."""

    def __init__(self, hidden, actionspace):
        super().__init__()
        self.linear = nn.Linear(hidden, actionspace)
        self.softmax = nn.LogSoftmax(dim=-1)


class BertPreTrainedModel(PreTrainedModel):
    """hf/modeling_bert.py:539-561."""
    base_model_prefix = "bert"


# ------------------------------------------------------------------------------------------------
# per-call state shared by the encoder node and the heads node
# ------------------------------------------------------------------------------------------------
class _Step:
    __slots__ = ("B", "T", "R", "S", "M", "seed", "p_hidden", "p_attn", "ids", "tt", "pos", "bias", "key_tiles", "has_img",
                 "img_cat", "img_pre", "layers", "seq", "pooled", "began", "model", "need_grad", "hidden", "plan", "rows", "cls_in")

    def __init__(self):
        self.layers = []
        self.img_cat = self.img_pre = None
        self.began = False
        self.hidden = None
        self.plan = None        # ops.PackPlan: the step runs on packed rows (padding-free); None = the dense [B, S] layout
        self.cls_in = None


# Set by visitron_b200.graph while a step is being captured: the dropout seed then lives in device memory (ops.set_seed_source)
# and is advanced by the first node of the graph; the by-value seed of every launch is a constant.
_CAPTURING = False


def _draw_seed() -> int:
    if _CAPTURING:
        return 0x5EED
    # CPU generator: reproducible under torch.manual_seed (vs/utils.py:24-29), no device sync
    s = int(torch.randint(0, 2 ** 62, (1,)).item())
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        s ^= (torch.distributed.get_rank() + 1) * 0x9E3779B97F4A7C15 & (2 ** 63 - 1)  # ranks must not share masks
    return s


# ------------------------------------------------------------------------------------------------
# the encoder model
# ------------------------------------------------------------------------------------------------
class BertImgModelwithLocationEmbeds(BertPreTrainedModel):
    """vs/encoder.py:161-303 -- expands BertModel to take image-region features as extra tokens."""

    def __init__(self, config):
        super().__init__(config)
        check_envelope(config)
        self.embeddings = BertEmbeddings(config)
        self.encoder = CaptionBertEncoder(config)
        self.pooler = BertPooler(config)
        self.img_dim = config.img_feature_dim
        logger.info("BertImgModel Image Dimension: {}".format(self.img_dim))
        self.img_feature_type = getattr(config, "img_feature_type", "faster_r-cnn")
        self.use_img_layernorm = getattr(config, "use_img_layernorm", None)
        self.img_embedding = nn.Linear(self.img_dim, config.hidden_size, bias=True)
        self.location_embeds = nn.Linear(LOC_DIM, config.hidden_size, bias=True)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        if self.use_img_layernorm:
            self.LayerNorm = BertLayerNorm(config.hidden_size, eps=config.img_layer_norm_eps)
        self.apply(self.init_weights)
        self._vt_store: Optional[FlatStore] = None
        self._vt_store_key = None
        self._vt_img_cache = None
        self._vt_reducer = None

    # -- reference API ---------------------------------------------------------------------------
    def resize_specific_embeddings(self, embedding_type, new_num_tokens):
        """vs/encoder.py:189-194; like the reference this does NOT re-tie the MLM decoder."""
        old = getattr(self.embeddings, embedding_type)
        new = self._get_resized_embeddings(old, new_num_tokens)
        setattr(self.embeddings, embedding_type, new)
        return getattr(self.embeddings, embedding_type)

    def _prune_heads(self, heads_to_prune):
        raise NotImplementedError("head pruning changes the 12 x 64 head geometry the CUDA kernels are specialised for")

    def _vt_set_reducer(self, reducer):
        self._vt_reducer = reducer

    # -- flat storage ----------------------------------------------------------------------------
    def _layout(self):
        e = self.embeddings
        named = [("embeddings.word_embeddings.weight", e.word_embeddings.weight),
                 ("embeddings.position_embeddings.weight", e.position_embeddings.weight),
                 ("embeddings.token_type_embeddings.weight", e.token_type_embeddings.weight),
                 ("embeddings.LayerNorm.weight", e.LayerNorm.weight), ("embeddings.LayerNorm.bias", e.LayerNorm.bias),
                 ("img_embedding.weight", self.img_embedding.weight), ("img_embedding.bias", self.img_embedding.bias),
                 ("location_embeds.weight", self.location_embeds.weight), ("location_embeds.bias", self.location_embeds.bias)]
        if self.use_img_layernorm:
            named += [("LayerNorm.weight", self.LayerNorm.weight), ("LayerNorm.bias", self.LayerNorm.bias)]
        starts = ["embeddings.word_embeddings.weight"]
        for i, L in enumerate(self.encoder.layer):
            p = "encoder.layer.%d." % i
            a, so = L.attention.self, L.attention.output
            named += [(p + "attention.self.query.weight", a.query.weight), (p + "attention.self.key.weight", a.key.weight),
                      (p + "attention.self.value.weight", a.value.weight), (p + "attention.self.query.bias", a.query.bias),
                      (p + "attention.self.key.bias", a.key.bias), (p + "attention.self.value.bias", a.value.bias),
                      (p + "attention.output.dense.weight", so.dense.weight), (p + "attention.output.dense.bias", so.dense.bias),
                      (p + "attention.output.LayerNorm.weight", so.LayerNorm.weight), (p + "attention.output.LayerNorm.bias", so.LayerNorm.bias),
                      (p + "intermediate.dense.weight", L.intermediate.dense.weight), (p + "intermediate.dense.bias", L.intermediate.dense.bias),
                      (p + "output.dense.weight", L.output.dense.weight), (p + "output.dense.bias", L.output.dense.bias),
                      (p + "output.LayerNorm.weight", L.output.LayerNorm.weight), (p + "output.LayerNorm.bias", L.output.LayerNorm.bias)]
            starts.append(p + "attention.self.query.weight")
        named += [("pooler.dense.weight", self.pooler.dense.weight), ("pooler.dense.bias", self.pooler.dense.bias)]
        return named, starts

    def _store(self) -> FlatStore:
        e = self.embeddings
        key = (id(e.word_embeddings.weight), id(e.position_embeddings.weight), id(e.token_type_embeddings.weight),
               tuple(e.word_embeddings.weight.shape), tuple(e.position_embeddings.weight.shape),
               tuple(e.token_type_embeddings.weight.shape))
        if self._vt_store is None or self._vt_store_key != key:
            named, starts = self._layout()
            self._vt_store = FlatStore(named, starts)
            self._vt_store_key = key
            self._vt_img_cache = None
        return self._vt_store.ensure()

    def _img_operands(self, st: FlatStore):
        """[W_loc | W_img | 0] as one bf16 [768, Kp] operand and b_img + b_loc: img_embedding(x) +
        location_embeds(l) (vs/encoder.py:277-279) is ONE contraction over the concatenated input
        (the 128 location columns come first so that both column blocks start 16-byte aligned)."""
        # re-cast at EVERY forward (two small launches + one add): like FlatStore.refresh_bf16, nothing the host can observe
        # proves that the masters did not change since the last call (in-place updates through p.data)
        key = (st.flat.data_ptr(), self.img_dim)
        if self._vt_img_cache is None or self._vt_img_cache[0] != key:
            kp = ops.roundup(self.img_dim + LOC_DIM, 8)
            w = torch.zeros((H, kp), dtype=BF16, device=st.flat.device)
            b = torch.empty(H, dtype=F32, device=st.flat.device)
            self._vt_img_cache = (key, w, b, kp)
        _, w, b, kp = self._vt_img_cache
        ops.cast_weight(st.w("location_embeds.weight"), w[:, :LOC_DIM])
        ops.cast_weight(st.w("img_embedding.weight"), w[:, LOC_DIM: LOC_DIM + self.img_dim])
        torch.add(st.w("img_embedding.bias"), st.w("location_embeds.bias"), out=b)
        return self._vt_img_cache[1:]

    # -- forward ---------------------------------------------------------------------------------
    def forward(self, input_ids, token_type_ids=None, attention_mask=None, position_ids=None, head_mask=None,
                img_feats=None, img_location_embeddings=None, encoder_history_states=None):
        seq, pooled, step = self._vt_forward(input_ids, token_type_ids, attention_mask, position_ids, head_mask,
                                             img_feats, img_location_embeddings, encoder_history_states)
        sequence_output = _ToF32.apply(seq).view(step.B, step.S, H)
        if step.hidden is not None:
            # config.output_hidden_states (oscar/modeling_bert.py:146-160): the embedding output and every layer's output, which the
            # fused path holds anyway; returned as fp32 COPIES outside the autograd graph (gradients flow through sequence_output)
            hs = tuple(ops.to_f32(h).view(step.B, step.S, H) for h in step.hidden[:-1]) + (sequence_output.detach(),)
            return (sequence_output, pooled, hs)
        return (sequence_output, pooled)

    def _vt_forward(self, input_ids, token_type_ids=None, attention_mask=None, position_ids=None, head_mask=None,
                    img_feats=None, img_location_embeddings=None, encoder_history_states=None, pack_labels=None):
        """Returns (sequence_output as bf16 [rows, 768], pooled_output fp32 [B, 768], step state).

        pack_labels = (labels, token_labels) flat int64 [B*S] asks for PADDING-FREE execution: the rows that are masked as keys and
        carry no label are dropped on the device (ops.pack_plan) and the encoder runs on the packed rows; step.plan then maps between
        the layouts and sequence_output is in packed order (only PreTrainOscar, whose heads read labelled rows, asks for it)."""
        cfg = self.config
        if head_mask is not None:
            raise NotImplementedError("head_mask is not supported by the fused attention kernel (no reference caller passes it)")
        if encoder_history_states:
            assert img_feats is None, "Cannot take image features while using encoder history states"
            raise NotImplementedError("encoder_history_states is not supported (no reference caller passes it)")
        if getattr(cfg, "output_attentions", False):
            raise NotImplementedError("output_attentions would materialise the S x S probabilities the fused attention kernel never "
                                      "writes (no reference caller sets it)")
        lib.require_device(input_ids)
        if input_ids.dim() != 2:
            raise ValueError("input_ids must be [batch, text_len]")
        B, T = input_ids.shape
        if T > self.embeddings.position_embeddings.weight.size(0) and position_ids is None:
            raise IndexError("text length %d exceeds the position table (%d)" % (T, self.embeddings.position_embeddings.weight.size(0)))
        step = _Step()
        step.model = self
        step.B, step.T = B, T
        step.has_img = img_feats is not None
        step.R = img_feats.size(1) if step.has_img else 0
        step.S = T + step.R
        step.M = B * step.S
        training = self.training
        step.p_hidden = float(cfg.hidden_dropout_prob) if training else 0.0
        step.p_attn = float(cfg.attention_probs_dropout_prob) if training else 0.0
        step.seed = _draw_seed() if (step.p_hidden > 0 or step.p_attn > 0) else 0
        step.ids = input_ids.contiguous().long()
        step.tt = None if token_type_ids is None else token_type_ids.expand(B, T).contiguous().long()
        step.pos = None if position_ids is None else position_ids.expand(B, T).contiguous().long()
        if attention_mask is None:
            attention_mask = torch.ones((B, step.S), dtype=torch.uint8, device=input_ids.device)  # ones_like(input_ids) in the reference
        if attention_mask.dim() == 3:
            raise NotImplementedError("3-D (per-query) attention masks are not supported by the fused attention kernel; "
                                      "no reference caller passes one (vs/encoder.py:228-229)")
        if attention_mask.dim() != 2:
            raise NotImplementedError
        if tuple(attention_mask.shape) != (B, step.S):
            raise ValueError("attention_mask must be [batch, text_len + regions] = %s, got %s" % ((B, step.S), tuple(attention_mask.shape)))
        step.bias = ops.mask_bias(attention_mask)
        step.rows = step.M
        if pack_labels is not None and not self.use_img_layernorm and not getattr(cfg, "output_hidden_states", False):
            step.plan = ops.pack_plan(step.bias, pack_labels[0], pack_labels[1], B, step.S)
            step.rows = step.plan.cap
            step.key_tiles = step.plan.key_tiles
        else:
            step.key_tiles = ops.attention_key_tiles(step.bias)   # once per step: the same padding for all layers, forward and backward
        if step.has_img:
            if img_location_embeddings is None:
                raise ValueError("img_location_embeddings is required with img_feats (vs/encoder.py:277-279)")
            if img_feats.size(2) != self.img_dim or img_location_embeddings.size(2) != LOC_DIM:
                raise ValueError("img_feats / img_location_embeddings have the wrong feature width")
        store = self._store()
        store.refresh_bf16()
        step.need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in store.params)
        if getattr(cfg, "output_hidden_states", False):
            step.hidden = []
        imgs = []
        if step.has_img:
            # fp32 as the loader delivers them, or bf16 from a feature cache (read as it is: half the H2D bytes, no fp32 detour)
            imgs = [t.contiguous() if t.dtype == BF16 else t.contiguous().float() for t in (img_feats, img_location_embeddings)]
        seq, pooled = _EncoderFn.apply(step, store, *imgs, *store.params)
        return seq, pooled, step


class _ToF32(torch.autograd.Function):
    """Module boundary: bf16 activations -> the fp32 tensors the reference returns."""

    @staticmethod
    def forward(ctx, x):
        return ops.to_f32(x)

    @staticmethod
    def backward(ctx, g):
        return ops.to_bf16(g.contiguous())


# ------------------------------------------------------------------------------------------------
# encoder node
# ------------------------------------------------------------------------------------------------
def _encoder_forward(step: _Step, st: FlatStore, img_feats, img_loc):
    model = step.model
    cfg = model.config
    B, T, R, S, M = step.B, step.T, step.R, step.S, step.M
    dev = step.ids.device
    eps = float(cfg.layer_norm_eps)
    nh = cfg.num_attention_heads
    I = cfg.intermediate_size
    save = step.need_grad
    plan = step.plan
    rows = step.rows                  # rows of every activation buffer: B*S, or the packed layout's capacity
    row_dst = None if plan is None else plan.row_dst
    # packed rows: the rows between the last sample's end and the row limit must be finite zeros (they ride along in every GEMM)
    x = torch.empty((rows, H), dtype=BF16, device=dev) if plan is None else torch.zeros((rows, H), dtype=BF16, device=dev)
    ops.embed_ln_fwd(step.ids, step.tt, step.pos, st.w("embeddings.word_embeddings.weight"),
                     st.w("embeddings.position_embeddings.weight"), st.w("embeddings.token_type_embeddings.weight"),
                     st.w("embeddings.LayerNorm.weight"), st.w("embeddings.LayerNorm.bias"), eps, x, S,
                     step.p_hidden, step.seed, STREAM_EMB, row_dst=row_dst)
    if step.has_img:
        wimg, bimg, kp = model._img_operands(st)
        cat = torch.empty((B * R, kp), dtype=BF16, device=dev)
        ops.pad_cast_concat(img_loc.view(B * R, -1), img_feats.view(B * R, -1), cat)
        if model.use_img_layernorm:
            pre = lib.gemm(cat, wimg, bias=bimg)
            ops.layernorm_fwd(pre, st.w("LayerNorm.weight"), st.w("LayerNorm.bias"), float(cfg.img_layer_norm_eps), out=x,
                              dropout_p=step.p_hidden, seed=step.seed, stream_id=STREAM_IMG, remap=(R, S, T))
            step.img_pre = pre if save else None
        else:
            lib.gemm(cat, wimg, bias=bimg, out=x, dropout_p=step.p_hidden, seed=step.seed, dropout_stream=STREAM_IMG,
                     remap=(R, S, T), remap_index=row_dst)
        step.img_cat = cat if save else None
    limit = ops.row_limit(plan.counts, 0) if plan is not None else contextlib.nullcontext()
    with limit:
        x = _encoder_layers_forward(step, st, x, save)
    if step.hidden is not None:
        step.hidden.append(x)
    if plan is None:
        cls_in = x[::S]
    else:
        cls_in = ops.gather_rows(x, plan.cls_rows)       # every sample's first row ([CLS]) in the packed layout
        step.cls_in = cls_in if save else None
    pooled = lib.gemm(cls_in, st.wb("pooler.dense.weight"), bias=st.w("pooler.dense.bias"), act=ACT_TANH, out_dtype=F32)
    # detached ALIASES: autograd attaches this node to the returned tensor objects themselves, and node -> ctx.step -> step.seq
    # -> grad_fn -> node would be a reference cycle that keeps every finished iteration's graph (and its AccumulateGrad nodes)
    # alive until the cyclic collector runs
    step.seq, step.pooled = x.detach(), pooled.detach()
    return x, pooled


def _encoder_layers_forward(step: _Step, st: FlatStore, x, save):
    """The 12 layers (oscar/modeling_bert.py:145-155); under a row limit when the step runs on packed rows."""
    cfg = step.model.config
    B, S, nh, I = step.B, step.S, cfg.num_attention_heads, cfg.intermediate_size
    M = step.rows
    dev = x.device
    eps = float(cfg.layer_norm_eps)
    for l in range(cfg.num_hidden_layers):
        p = "encoder.layer.%d." % l
        if step.hidden is not None:
            step.hidden.append(x)
        qkv = lib.gemm(x, st.wb(p + "attention.self.query.weight", 3 * H, H), bias=st.wspan(p + "attention.self.query.bias", 3 * H))
        # the probability dropout's keep decisions are stored (1 bit each) for the backward, which then does not draw them again
        bits = ops.attention_keep_bits(B, S, nh, dev) if (save and step.p_attn > 0) else None
        ctx, lse = ops.attention_fwd(qkv, step.bias, B, S, nh, step.p_attn, step.seed, 4 * l, key_tiles=step.key_tiles, plan=step.plan,
                                     keep_bits=bits)
        x1 = lib.gemm(ctx, st.wb(p + "attention.output.dense.weight"), bias=st.w(p + "attention.output.dense.bias"),
                      dropout_p=step.p_hidden, seed=step.seed, dropout_stream=4 * l + 1, residual=x)
        y1 = ops.layernorm_fwd(x1, st.w(p + "attention.output.LayerNorm.weight"), st.w(p + "attention.output.LayerNorm.bias"), eps)
        # when the backward will run, the FFN-up epilogue saves gelu'(pre) (from the fp32 accumulator) instead of pre: the backward GEMM's
        # epilogue is then one multiply
        pre = torch.empty((M, I), dtype=BF16, device=dev) if save else None
        z = lib.gemm(y1, st.wb(p + "intermediate.dense.weight"), bias=st.w(p + "intermediate.dense.bias"),
                     act=ACT_GELU_SAVE_GRAD if save else ACT_GELU, aux_out=pre)
        x2 = lib.gemm(z, st.wb(p + "output.dense.weight"), bias=st.w(p + "output.dense.bias"),
                      dropout_p=step.p_hidden, seed=step.seed, dropout_stream=4 * l + 2, residual=y1)
        y2 = ops.layernorm_fwd(x2, st.w(p + "output.LayerNorm.weight"), st.w(p + "output.LayerNorm.bias"), eps)
        if save:
            step.layers.append((x, qkv, ctx, lse, x1, y1, pre, z, x2, bits))
        x = y2
    return x


def _begin_backward(step: _Step, stores: List[FlatStore]):
    if step.began:
        return
    step.began = True
    for s in stores:
        s.begin_backward()


def _encoder_backward(step: _Step, st: FlatStore, d, dpooled):
    """d: bf16 [M, 768] gradient of the sequence output (modified in place) or None."""
    model = step.model
    cfg = model.config
    B, T, R, S, M = step.B, step.T, step.R, step.S, step.M
    eps = float(cfg.layer_norm_eps)
    nh = cfg.num_attention_heads
    red = model._vt_reducer
    plan = step.plan
    row_dst = None if plan is None else plan.row_dst
    if d is None:
        d = torch.zeros((step.rows, H), dtype=BF16, device=step.ids.device)
    x_last = step.seq
    if dpooled is not None:
        dp = ops.tanh_bwd(dpooled.contiguous(), step.pooled)
        cls_in = x_last[::S] if plan is None else step.cls_in
        lib.gemm(dp, cls_in, a_major=1, b_major=1, out=st.g("pooler.dense.weight"), accumulate=True, colsum=st.g("pooler.dense.bias"))
        lib.gemm(dp, st.wb("pooler.dense.weight"), b_major=1, out=d, residual=d, remap=(1, S, 0), remap_index=row_dst)
    limit = ops.row_limit(plan.counts, 0) if plan is not None else contextlib.nullcontext()
    with limit:
        d = _encoder_layers_backward(step, st, d)
    if step.has_img:
        cat = step.img_cat
        if model.use_img_layernorm:
            dimg, _ = ops.layernorm_bwd(d, step.img_pre, st.w("LayerNorm.weight"), float(cfg.img_layer_norm_eps),
                                        st.g("LayerNorm.weight"), st.g("LayerNorm.bias"), st.g("img_embedding.bias"),
                                        in_dropout_p=step.p_hidden, in_stream=STREAM_IMG, seed=step.seed, remap=(R, S, T))
            st.g("location_embeds.bias").copy_(st.g("img_embedding.bias"))
        else:
            dimg = torch.empty((B * R, H), dtype=BF16, device=d.device)
            ops.dropout_rows(d, B * R, dimg, step.p_hidden, step.seed, STREAM_IMG, remap=(R, S, T), row_dst=row_dst)
            ops.colsum(dimg, st.g("img_embedding.bias"))
            ops.colsum(dimg, st.g("location_embeds.bias"))
        D = model.img_dim
        lib.gemm(dimg, cat[:, LOC_DIM: LOC_DIM + D], a_major=1, b_major=1, out=st.g("img_embedding.weight"), accumulate=True)
        lib.gemm(dimg, cat[:, :LOC_DIM], a_major=1, b_major=1, out=st.g("location_embeds.weight"), accumulate=True)
    ops.embed_ln_bwd(d, step.ids, step.tt, step.pos, st.w("embeddings.word_embeddings.weight"),
                     st.w("embeddings.position_embeddings.weight"), st.w("embeddings.token_type_embeddings.weight"),
                     st.w("embeddings.LayerNorm.weight"), eps, S, st.g("embeddings.word_embeddings.weight"),
                     st.g("embeddings.position_embeddings.weight"), st.g("embeddings.token_type_embeddings.weight"),
                     st.g("embeddings.LayerNorm.weight"), st.g("embeddings.LayerNorm.bias"), step.p_hidden, step.seed, STREAM_EMB,
                     row_dst=row_dst)
    if red is not None:
        red.reduce(st.grad_bucket(0), tail=True)  # embeddings (+ the tied MLM decoder's contribution) and the image projection
        red.finish()


def _encoder_layers_backward(step: _Step, st: FlatStore, d):
    """Backward of the 12 layers; under a row limit when the step runs on packed rows."""
    model = step.model
    cfg = model.config
    B, S = step.B, step.S
    eps = float(cfg.layer_norm_eps)
    nh = cfg.num_attention_heads
    red = model._vt_reducer
    for l in reversed(range(cfg.num_hidden_layers)):
        p = "encoder.layer.%d." % l
        x_in, qkv, ctx, lse, x1, y1, pre, z, x2, bits = step.layers[l]
        step.layers[l] = None
        ph = step.p_hidden
        # LN2 and the FFN
        dx2, dx2d = ops.layernorm_bwd(d, x2, st.w(p + "output.LayerNorm.weight"), eps, st.g(p + "output.LayerNorm.weight"),
                                      st.g(p + "output.LayerNorm.bias"), st.g(p + "output.dense.bias"),
                                      out_dropout_p=ph, out_stream=4 * l + 2, seed=step.seed)
        # (these two bias gradients come from the LayerNorm backward's third accumulator: without the column-sum side product the
        # weight-gradient GEMM can run on CTA pairs, 109 -> 87 us for dW[768, 3072] at 24.5 k rows)
        lib.gemm(dx2d, z, a_major=1, b_major=1, out=st.g(p + "output.dense.weight"), accumulate=True)
        dpre = lib.gemm(dx2d, st.wb(p + "output.dense.weight"), b_major=1, act=ACT_MUL_AUX, aux=pre)   # pre holds gelu'(pre-activation)
        # the bias gradients ride on the weight-gradient GEMMs (column sums of the dY tiles they stage anyway)
        lib.gemm(dpre, y1, a_major=1, b_major=1, out=st.g(p + "intermediate.dense.weight"), accumulate=True,
                 colsum=st.g(p + "intermediate.dense.bias"))
        dy1 = lib.gemm(dpre, st.wb(p + "intermediate.dense.weight"), b_major=1, residual=dx2)
        del dpre, dx2, dx2d, pre, z, x2
        # LN1 and the attention block
        dx1, dx1d = ops.layernorm_bwd(dy1, x1, st.w(p + "attention.output.LayerNorm.weight"), eps,
                                      st.g(p + "attention.output.LayerNorm.weight"), st.g(p + "attention.output.LayerNorm.bias"),
                                      st.g(p + "attention.output.dense.bias"), out_dropout_p=ph, out_stream=4 * l + 1, seed=step.seed)
        lib.gemm(dx1d, ctx, a_major=1, b_major=1, out=st.g(p + "attention.output.dense.weight"), accumulate=True)
        dctx = lib.gemm(dx1d, st.wb(p + "attention.output.dense.weight"), b_major=1)
        dqkv = ops.attention_bwd(qkv, step.bias, ctx, dctx, lse, B, S, nh, step.p_attn, step.seed, 4 * l, key_tiles=step.key_tiles, plan=step.plan,
                                 keep_bits=bits)
        lib.gemm(dqkv, x_in, a_major=1, b_major=1, out=st.g(p + "attention.self.query.weight", 3 * H, H), accumulate=True,
                 colsum=st.g(p + "attention.self.query.bias", 3 * H))
        d = lib.gemm(dqkv, st.wb(p + "attention.self.query.weight", 3 * H, H), b_major=1, residual=dx1)
        del dqkv, dctx, dx1, dx1d, dy1, qkv, ctx, x1, y1, x_in
        if red is not None:  # this layer's slice (the last layer's also holds the pooler) is complete: exchange it now
            red.reduce(st.grad_bucket(st.bucket_of(p + "attention.self.query.weight")))
    return d


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, step: _Step, store: FlatStore, *tensors):
        n_img = 2 if step.has_img else 0
        img_feats, img_loc = (tensors[0], tensors[1]) if n_img else (None, None)
        seq, pooled = _encoder_forward(step, store, img_feats, img_loc)
        ctx.step, ctx.store, ctx.n_img = step, store, n_img
        ctx.set_materialize_grads(False)  # an unused pooled_output (fine-tuning) must arrive as None, not zeros
        return seq, pooled

    @staticmethod
    def backward(ctx, dseq, dpooled):
        step, store = ctx.step, ctx.store
        if step.layers is None or (len(step.layers) and step.layers[-1] is None):
            raise RuntimeError("the visitron_b200 encoder frees its activations during backward; backward through the same "
                               "forward twice (retain_graph) is not supported")
        _begin_backward(step, [store])
        if dseq is not None and not dseq.is_contiguous():
            dseq = dseq.contiguous()
        _encoder_backward(step, store, dseq, dpooled)
        used = None
        if not step.has_img:
            used = set(store.names) - {"img_embedding.weight", "img_embedding.bias", "location_embeds.weight",
                                       "location_embeds.bias", "LayerNorm.weight", "LayerNorm.bias"}
        if dpooled is None:
            used = (set(store.names) if used is None else used) - {"pooler.dense.weight", "pooler.dense.bias"}
        step.layers = None
        if _CAPTURING:
            store.assign_grads(used)
            return (None, None) + (None,) * ctx.n_img + (None,) * len(store.params)
        return (None, None) + (None,) * ctx.n_img + tuple(store.grad_views(used))


# ------------------------------------------------------------------------------------------------
# pre-training model
# ------------------------------------------------------------------------------------------------
class PreTrainOscar(BertPreTrainedModel):
    """vs/encoder.py:306-441: encoder + masked-LM head (tied decoder) + masked detector-class head
    (Softmax followed by CrossEntropyLoss, i.e. a double softmax) + 36-way next-action head."""

    def __init__(self, config):
        super().__init__(config)
        self.config = config
        self.bert = BertImgModelwithLocationEmbeds(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.next_action = NextActionPrediction(config.hidden_size, config.action_space)
        self.criterion = nn.CrossEntropyLoss(ignore_index=-1)
        self.mlmhead = BertOnlyMLMHead(config)
        self.token_head = nn.Sequential(nn.Linear(config.hidden_size, config.detector_classes), nn.Softmax(dim=-1))
        self.apply(self.init_weights)
        self.tie_weights()
        self._vt_store: Optional[FlatStore] = None
        self._vt_store_key = None
        self._vt_reducer = None
        self._vt_aux_raw = None
        self.vt_packed = True      # run the step on the rows that exist (visitron_b200/csrc/pack.cu); False: the dense [B, S] layout

    def tie_weights(self):
        self._tie_or_clone_weights(self.mlmhead.predictions.decoder, self.bert.embeddings.word_embeddings)

    def resize_embeddings(self, embedding_size_dict):
        for embedding_type, new_embedding_size in embedding_size_dict.items():
            assert embedding_type in ["word_embeddings", "position_embeddings", "token_type_embeddings"]
            self.bert.resize_specific_embeddings(embedding_type, new_embedding_size)
            logger.info(f"Resized {embedding_type} to {new_embedding_size}")

    def _vt_set_reducer(self, reducer):
        self._vt_reducer = reducer

    def _store(self) -> FlatStore:
        dec = self.mlmhead.predictions.decoder.weight
        tied = dec is self.bert.embeddings.word_embeddings.weight
        key = (tied, id(dec), tuple(dec.shape))
        if self._vt_store is None or self._vt_store_key != key:
            t = self.mlmhead.predictions.transform
            named = [("mlmhead.predictions.transform.dense.weight", t.dense.weight), ("mlmhead.predictions.transform.dense.bias", t.dense.bias),
                     ("mlmhead.predictions.transform.LayerNorm.weight", t.LayerNorm.weight),
                     ("mlmhead.predictions.transform.LayerNorm.bias", t.LayerNorm.bias),
                     ("mlmhead.predictions.bias", self.mlmhead.predictions.bias),
                     ("token_head.0.weight", self.token_head[0].weight), ("token_head.0.bias", self.token_head[0].bias),
                     ("next_action.linear.weight", self.next_action.linear.weight), ("next_action.linear.bias", self.next_action.linear.bias)]
            if not tied:
                # after resize_embeddings (vs/model_utils.py:101-109, `--no_oscar_setting`) the decoder keeps pointing at the OLD
                # [30522, 768] table (vs/encoder.py:337-345 does not re-tie): it is then a parameter of its own, trained as one
                named.append(("mlmhead.predictions.decoder.weight", dec))
            self._vt_store = FlatStore(named)
            self._vt_store_key = key
        return self._vt_store.ensure()

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, token_labels=None, position_ids=None,
                head_mask=None, img_feats=None, img_location_embeddings=None, next_action=None, text_only=False):
        if text_only:
            return self.bert(input_ids, position_ids=position_ids, token_type_ids=token_type_ids, attention_mask=attention_mask,
                             head_mask=head_mask, img_feats=img_feats, img_location_embeddings=img_location_embeddings)
        if labels is None or token_labels is None:
            # the reference itself fails here: labels.view(-1) on None, resp. UnboundLocalError on
            # token_prediction (vs/encoder.py:380,388,400)
            raise ValueError("PreTrainOscar.forward needs `labels` and `token_labels` (the reference crashes without them)")
        lib.require_device(input_ids)
        dec = self.mlmhead.predictions.decoder.weight
        V = dec.size(0)
        if self.mlmhead.predictions.bias.numel() != V:
            raise ValueError("mlmhead.predictions.bias has %d entries for a decoder of %d rows" % (self.mlmhead.predictions.bias.numel(), V))
        lab = labels.reshape(-1).contiguous().long()
        tok = token_labels.reshape(-1).contiguous().long()
        S_all = input_ids.size(1) + (img_feats.size(1) if img_feats is not None else 0)
        if lab.numel() != input_ids.size(0) * S_all or tok.numel() != lab.numel():
            raise ValueError("labels / token_labels must be [batch, text_len + regions]")
        # padding-free execution (self.vt_packed, default on): the rows that are masked out and unlabelled are dropped on the device
        seq, pooled, step = self.bert._vt_forward(input_ids, token_type_ids, attention_mask, position_ids, head_mask,
                                                  img_feats, img_location_embeddings, pack_labels=(lab, tok) if self.vt_packed else None)
        if step.plan is not None:
            lab, tok = step.plan.labels_a, step.plan.labels_b      # the labels in packed row order (-1 on the layout's padding rows)
        hs = self._store()
        hs.refresh_bf16()
        na = None if next_action is None else next_action.contiguous().long()
        # the heads save their activations whenever ANY gradient can flow: a frozen encoder with trainable heads included
        save = torch.is_grad_enabled() and (step.need_grad or any(p.requires_grad for p in hs.params))
        out8 = _HeadsFn.apply(self, step, hs, (lab, tok, na, V, save), seq, pooled, *hs.params)
        return (out8[0], out8[1], out8[2], out8[3], out8[4], out8[5], out8[6])

    @property
    def _vt_last_aux(self):
        """Head outputs of the last forward, cut to the labelled rows (test / debugging accessor: reads the row counts, i.e. SYNCS)."""
        raw = self._vt_aux_raw
        c = raw["counts"].cpu()
        n, nt = int(c[1]), int(c[5])
        if raw.get("row_src") is not None:      # packed rows: report the labelled rows in the caller's [B, S] numbering
            raw = dict(raw, idx_mlm=raw["row_src"][raw["idx_mlm"][:n]].long(), idx_tok=raw["row_src"][raw["idx_tok"][:nt]].long())
        return dict(predicted_words=raw["predicted_words"][:n], predicted_tokens=raw["predicted_tokens"][:nt],
                    predicted_action=raw["predicted_action"], action_scores=raw["action_scores"],
                    mlm_logits=raw["mlm_logits"][:n], token_logits=raw["token_logits"][:nt],
                    mlm_rows=raw["idx_mlm"][:n], token_rows=raw["idx_tok"][:nt], label_errors=int(c[2]) + int(c[6]))


class _HeadsFn(torch.autograd.Function):
    """The three pre-training heads and the loss assembly (vs/encoder.py:377-441) on the LABELLED rows only.  Which rows those
    are is decided on the device (ops.compact_labels); every buffer is sized for the bound (all M rows, rounded up to 128)
    and every launch clips itself to the device-resident count (ops.row_limit), so the step has no host <-> device sync and a
    fixed launch sequence (CUDA-graph capturable)."""

    @staticmethod
    def forward(ctx, model: PreTrainOscar, step: _Step, hs: FlatStore, meta, seq, pooled, *params):
        lab, tok, next_action, V, save = meta
        cfg = model.config
        bs: FlatStore = step.model._vt_store
        eps = float(cfg.layer_norm_eps)
        Dc = cfg.detector_classes
        dev = seq.device
        cap = ops.roundup(lab.numel(), 128)
        idx_mlm, lab_mlm, idx_tok, lab_tok, counts = ops.compact_labels(lab, V, tok, Dc, cap)
        tied = "mlmhead.predictions.decoder.weight" not in hs.offset
        E = bs.wb("embeddings.word_embeddings.weight") if tied else hs.wb("mlmhead.predictions.decoder.weight")
        # masked-LM head on the labelled rows (the only rows that reach loss, accuracy or gradients)
        ldv = ops.roundup(V, 8)
        logits = torch.empty((cap, ldv), dtype=F32, device=dev)
        pre_t = torch.empty((cap, H), dtype=BF16, device=dev)
        with ops.row_limit(counts, 0):
            hm = ops.gather_rows(seq, idx_mlm)
            a = lib.gemm(hm, hs.wb("mlmhead.predictions.transform.dense.weight"), bias=hs.w("mlmhead.predictions.transform.dense.bias"),
                         act=ACT_GELU, aux_out=pre_t)
            tl = ops.layernorm_fwd(a, hs.w("mlmhead.predictions.transform.LayerNorm.weight"),
                                   hs.w("mlmhead.predictions.transform.LayerNorm.bias"), eps)
            lib.gemm(tl, E, bias=hs.w("mlmhead.predictions.bias"), out=logits[:, :V])
            mlm_loss, mlm_arg, mlm_stats = ops.softmax_ce_fwd(logits[:, :V], V, lab_mlm)
        # detector-class head (double softmax)
        ldt = ops.roundup(Dc, 8)
        tlogits = torch.empty((cap, ldt), dtype=F32, device=dev)
        with ops.row_limit(counts, 1):
            ht = ops.gather_rows(seq, idx_tok)
            lib.gemm(ht, hs.wb("token_head.0.weight"), bias=hs.w("token_head.0.bias"), out=tlogits[:, :Dc])
            tok_loss, tok_arg, tok_stats = ops.softmax_ce_fwd(tlogits[:, :Dc], Dc, lab_tok, double_softmax=True)
        # next-action head
        scores, act_loss, act_arg = ops.action_head_fwd(pooled, hs.w("next_action.linear.weight"), hs.w("next_action.linear.bias"),
                                                        next_action)
        out8 = ops.pretrain_finalize_counted((mlm_loss, mlm_arg, lab_mlm), (tok_loss, tok_arg, lab_tok),
                                             (act_loss, act_arg, next_action) if next_action is not None else (act_loss[:0], act_arg[:0], None),
                                             counts)
        ctx.model, ctx.step, ctx.hs = model, step, hs
        ctx.meta = (idx_mlm, lab_mlm, idx_tok, lab_tok, next_action, V, counts, tied)
        # out8.detach(): the returned out8 gets this node as grad_fn; saving the object itself would be a reference cycle
        ctx.saved = (hm, pre_t, a, tl, logits, mlm_stats, ht, tlogits, tok_stats, scores, out8.detach(), pooled) if save else None
        ctx.shapes = (cap, ldv, ldt)
        model._vt_aux_raw = dict(predicted_words=mlm_arg, predicted_tokens=tok_arg, predicted_action=act_arg, action_scores=scores,
                                 mlm_logits=logits[:, :V], token_logits=tlogits[:, :Dc], counts=counts, idx_mlm=idx_mlm, idx_tok=idx_tok,
                                 row_src=None if step.plan is None else step.plan.row_src)
        return out8

    @staticmethod
    def backward(ctx, g8):
        model, step, hs = ctx.model, ctx.step, ctx.hs
        idx_mlm, lab_mlm, idx_tok, lab_tok, next_action, V, counts, tied = ctx.meta
        if ctx.saved is None:
            raise RuntimeError("PreTrainOscar: backward through a forward that ran without gradients enabled")
        hm, pre_t, a, tl, logits, mlm_stats, ht, tlogits, tok_stats, scores, out8, pooled = ctx.saved
        ctx.saved = None
        cap, ldv, ldt = ctx.shapes
        cfg = model.config
        bs: FlatStore = step.model._vt_store
        eps = float(cfg.layer_norm_eps)
        Dc = cfg.detector_classes
        _begin_backward(step, [bs, hs])
        g8 = g8.contiguous().float()
        g_mlm = (g8[0:1] + g8[1:2]).contiguous()
        g_next = (g8[0:1] + g8[2:3]).contiguous()
        g_tok = (g8[0:1] + g8[3:4]).contiguous()
        dseq = torch.zeros((step.rows, H), dtype=BF16, device=hm.device)
        E = bs.wb("embeddings.word_embeddings.weight") if tied else hs.wb("mlmhead.predictions.decoder.weight")
        dE = bs.g("embeddings.word_embeddings.weight") if tied else hs.g("mlmhead.predictions.decoder.weight")
        with ops.row_limit(counts, 0):
            # the mean's 1 / n is applied on the device (the count lives there); padding rows of dlog are zeros
            dlog = ops.softmax_ce_bwd(logits[:, :V], V, lab_mlm, mlm_stats, 1.0, g_mlm, ld_out=ldv)
            lib.gemm(dlog[:, :V], tl, a_major=1, b_major=1, out=dE, accumulate=True, colsum=hs.g("mlmhead.predictions.bias"))
            dtl = lib.gemm(dlog[:, :V], E, b_major=1)
            del dlog
            da, _ = ops.layernorm_bwd(dtl, a, hs.w("mlmhead.predictions.transform.LayerNorm.weight"), eps,
                                      hs.g("mlmhead.predictions.transform.LayerNorm.weight"),
                                      hs.g("mlmhead.predictions.transform.LayerNorm.bias"), None)
            dpre = ops.gelu_bwd(da, pre_t)
            lib.gemm(dpre, hm, a_major=1, b_major=1, out=hs.g("mlmhead.predictions.transform.dense.weight"), accumulate=True,
                     colsum=hs.g("mlmhead.predictions.transform.dense.bias"))
            dhm = lib.gemm(dpre, hs.wb("mlmhead.predictions.transform.dense.weight"), b_major=1)
            ops.scatter_add_rows(dhm, idx_mlm, dseq)
        with ops.row_limit(counts, 1):
            dlt = ops.softmax_ce_bwd(tlogits[:, :Dc], Dc, lab_tok, tok_stats, 1.0, g_tok, double_softmax=True, ld_out=ldt)
            lib.gemm(dlt[:, :Dc], ht, a_major=1, b_major=1, out=hs.g("token_head.0.weight"), accumulate=True, colsum=hs.g("token_head.0.bias"))
            dht = lib.gemm(dlt[:, :Dc], hs.wb("token_head.0.weight"), b_major=1)
            ops.scatter_add_rows(dht, idx_tok, dseq)
        dpooled = None
        if next_action is not None:
            dpooled = ops.action_head_bwd(pooled, hs.w("next_action.linear.weight"), next_action, scores, out8[7:8], g_next,
                                          hs.g("next_action.linear.weight"), hs.g("next_action.linear.bias"))
        red = model._vt_reducer
        if red is not None:
            red.reduce(hs.grad_bucket(0))
        used = set(hs.names)
        if next_action is None:
            used -= {"next_action.linear.weight", "next_action.linear.bias"}
        if _CAPTURING:
            hs.assign_grads(used)
            return (None, None, None, None, dseq, dpooled) + (None,) * len(hs.params)
        return (None, None, None, None, dseq, dpooled) + tuple(hs.grad_views(used))
