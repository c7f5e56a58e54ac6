"""CPU oracle for the VISITRON encoder hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

This file restates, as plain functional fp32 PyTorch over a flat ``{state_dict key: tensor}``
dictionary, the arithmetic of the reference modules on the hot path.  It is
the checker for the CUDA path; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  Nothing under ``visitron_b200/`` does.

Pinning: the reference ships NO golden vectors for this path, so the
restatement is pinned against the *reference modules themselves* run in the build container
(``oracle/ref_shim.py`` imports them from /root/reference; ``oracle/make_golden.py`` dumps their
outputs on seeded inputs to ``tests/golden/*.pt``; ``tests/test_oracle.py`` compares).  On the GPU
box /root/reference does not exist; there the committed fixtures are the pin.

Each function cites the reference file:line it follows (paths relative to /root/reference;
``hf/`` = ``transformers/pytorch_transformers/``, ``vs/`` = ``tasks/viewpoint_select/``,
``tb/`` = ``tasks/turn_based/``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch
import torch.nn.functional as F

Params = Dict[str, torch.Tensor]


@dataclass
class OracleConfig:
    """The subset of BertConfig the path reads."""

    vocab_size: int = 30522
    hidden_size: int = 768
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    layer_norm_eps: float = 1e-12
    initializer_range: float = 0.02
    img_feature_dim: int = 2054
    loc_dim: int = 128  # vs/encoder.py:179 hard-codes 128
    use_img_layernorm: int = 0
    img_layer_norm_eps: float = 1e-12
    action_space: int = 36
    detector_classes: int = 1601
    hidden_dropout_prob: float = 0.0
    attention_probs_dropout_prob: float = 0.0


# --------------------------------------------------------------------------------------------
# parameters
# --------------------------------------------------------------------------------------------
def param_shapes(cfg: OracleConfig, pretrain_heads: bool = True) -> Dict[str, tuple]:
    """state_dict keys/shapes of PreTrainOscar (vs/encoder.py:306-345) in registration order."""
    H, I = cfg.hidden_size, cfg.intermediate_size
    s: Dict[str, tuple] = {}
    s["bert.embeddings.word_embeddings.weight"] = (cfg.vocab_size, H)
    s["bert.embeddings.position_embeddings.weight"] = (cfg.max_position_embeddings, H)
    s["bert.embeddings.token_type_embeddings.weight"] = (cfg.type_vocab_size, H)
    s["bert.embeddings.LayerNorm.weight"] = (H,)
    s["bert.embeddings.LayerNorm.bias"] = (H,)
    for i in range(cfg.num_hidden_layers):
        p = f"bert.encoder.layer.{i}."
        for n in ("query", "key", "value"):
            s[p + f"attention.self.{n}.weight"] = (H, H)
            s[p + f"attention.self.{n}.bias"] = (H,)
        s[p + "attention.output.dense.weight"] = (H, H)
        s[p + "attention.output.dense.bias"] = (H,)
        s[p + "attention.output.LayerNorm.weight"] = (H,)
        s[p + "attention.output.LayerNorm.bias"] = (H,)
        s[p + "intermediate.dense.weight"] = (I, H)
        s[p + "intermediate.dense.bias"] = (I,)
        s[p + "output.dense.weight"] = (H, I)
        s[p + "output.dense.bias"] = (H,)
        s[p + "output.LayerNorm.weight"] = (H,)
        s[p + "output.LayerNorm.bias"] = (H,)
    s["bert.pooler.dense.weight"] = (H, H)
    s["bert.pooler.dense.bias"] = (H,)
    s["bert.img_embedding.weight"] = (H, cfg.img_feature_dim)
    s["bert.img_embedding.bias"] = (H,)
    s["bert.location_embeds.weight"] = (H, cfg.loc_dim)
    s["bert.location_embeds.bias"] = (H,)
    if cfg.use_img_layernorm:
        s["bert.LayerNorm.weight"] = (H,)
        s["bert.LayerNorm.bias"] = (H,)
    if pretrain_heads:
        s["next_action.linear.weight"] = (cfg.action_space, H)
        s["next_action.linear.bias"] = (cfg.action_space,)
        s["mlmhead.predictions.bias"] = (cfg.vocab_size,)
        s["mlmhead.predictions.transform.dense.weight"] = (H, H)
        s["mlmhead.predictions.transform.dense.bias"] = (H,)
        s["mlmhead.predictions.transform.LayerNorm.weight"] = (H,)
        s["mlmhead.predictions.transform.LayerNorm.bias"] = (H,)
        # mlmhead.predictions.decoder.weight is TIED to the word embeddings (vs/encoder.py:332-335)
        s["token_head.0.weight"] = (cfg.detector_classes, H)
        s["token_head.0.bias"] = (cfg.detector_classes,)
    return s


def init_params(cfg: OracleConfig, seed: int, pretrain_heads: bool = True,
                dtype=torch.float32, perturb: bool = True) -> Params:
    """Seeded parameters.  hf/modeling_bert.py:550-561 draws Linear/Embedding weights from
    N(0, initializer_range), LayerNorm weight 1 / bias 0, Linear bias 0.  With ``perturb`` the
    biases and LayerNorm affine terms get small seeded noise on top, so that parity tests
    exercise every bias / gamma / beta path (a fresh init would multiply them away)."""
    g = torch.Generator().manual_seed(seed)
    out: Params = {}
    for k, shp in param_shapes(cfg, pretrain_heads).items():
        if k.endswith("LayerNorm.weight"):
            t = torch.ones(shp)
            if perturb:
                t += 0.1 * torch.randn(shp, generator=g)
        elif k.endswith("bias"):
            t = torch.zeros(shp)
            if perturb:
                t += 0.02 * torch.randn(shp, generator=g)
        else:
            t = cfg.initializer_range * torch.randn(shp, generator=g)
        out[k] = t.to(dtype)
    return out


# --------------------------------------------------------------------------------------------
# leaf blocks
# --------------------------------------------------------------------------------------------
def bert_layer_norm(x, w, b, eps):
    """hf/modeling_bert.py:227-240 -- TF style: biased variance, eps inside the sqrt."""
    u = x.mean(-1, keepdim=True)
    s = (x - u).pow(2).mean(-1, keepdim=True)
    return w * ((x - u) / torch.sqrt(s + eps)) + b


def gelu_erf(x):
    """hf/modeling_bert.py:136-142."""
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def extended_mask(attention_mask, dtype):
    """vs/encoder.py:226-241: (1 - mask.to(param dtype)) * -10000 for ANY integer mask value
    (fine-tune passes ~uint8 == 255/254)."""
    if attention_mask.dim() == 2:
        ext = attention_mask[:, None, None, :]
    elif attention_mask.dim() == 3:
        ext = attention_mask[:, None, :, :]
    else:
        raise NotImplementedError
    return (1.0 - ext.to(dtype)) * -10000.0


def embeddings(P: Params, cfg, input_ids, token_type_ids=None, position_ids=None, drop=None):
    """hf/modeling_bert.py:256-271."""
    T = input_ids.size(1)
    if position_ids is None:
        position_ids = torch.arange(T, dtype=torch.long, device=input_ids.device)[None].expand_as(input_ids)
    if token_type_ids is None:
        token_type_ids = torch.zeros_like(input_ids)
    e = (P["bert.embeddings.word_embeddings.weight"][input_ids]
         + P["bert.embeddings.position_embeddings.weight"][position_ids]
         + P["bert.embeddings.token_type_embeddings.weight"][token_type_ids])
    e = bert_layer_norm(e, P["bert.embeddings.LayerNorm.weight"], P["bert.embeddings.LayerNorm.bias"],
                        cfg.layer_norm_eps)
    return drop(e, cfg.hidden_dropout_prob) if drop else e


def self_attention(P: Params, cfg, pre: str, x, ext_mask, drop=None, return_probs=False):
    """oscar/modeling_bert.py:34-79 (history_state=None, head_mask=None)."""
    B, S, H = x.shape
    nh = cfg.num_attention_heads
    d = H // nh

    def heads(t):
        return t.view(B, S, nh, d).permute(0, 2, 1, 3)

    q = heads(F.linear(x, P[pre + "query.weight"], P[pre + "query.bias"]))
    k = heads(F.linear(x, P[pre + "key.weight"], P[pre + "key.bias"]))
    v = heads(F.linear(x, P[pre + "value.weight"], P[pre + "value.bias"]))
    scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(d)
    scores = scores + ext_mask
    probs = torch.softmax(scores, dim=-1)
    if drop:
        probs = drop(probs, cfg.attention_probs_dropout_prob)
    ctx = torch.matmul(probs, v).permute(0, 2, 1, 3).contiguous().view(B, S, H)
    return (ctx, probs) if return_probs else ctx


def encoder_layer(P: Params, cfg, i: int, x, ext_mask, drop=None):
    """oscar/modeling_bert.py:112-124 -> hf/modeling_bert.py:341-345, 387-390, 400-404."""
    p = f"bert.encoder.layer.{i}."
    ctx = self_attention(P, cfg, p + "attention.self.", x, ext_mask, drop)
    h = F.linear(ctx, P[p + "attention.output.dense.weight"], P[p + "attention.output.dense.bias"])
    if drop:
        h = drop(h, cfg.hidden_dropout_prob)
    y = bert_layer_norm(h + x, P[p + "attention.output.LayerNorm.weight"],
                        P[p + "attention.output.LayerNorm.bias"], cfg.layer_norm_eps)
    z = gelu_erf(F.linear(y, P[p + "intermediate.dense.weight"], P[p + "intermediate.dense.bias"]))
    o = F.linear(z, P[p + "output.dense.weight"], P[p + "output.dense.bias"])
    if drop:
        o = drop(o, cfg.hidden_dropout_prob)
    return bert_layer_norm(o + y, P[p + "output.LayerNorm.weight"], P[p + "output.LayerNorm.bias"],
                           cfg.layer_norm_eps)


def bert_img_forward(P: Params, cfg: OracleConfig, input_ids, token_type_ids=None, attention_mask=None,
                     position_ids=None, img_feats=None, img_location_embeddings=None, drop=None,
                     return_hidden=False):
    """BertImgModelwithLocationEmbeds.forward, vs/encoder.py:204-303 (head_mask=None,
    encoder_history_states=None).  Returns (sequence_output [B,S,H], pooled_output [B,H])."""
    if attention_mask is None:
        attention_mask = torch.ones_like(input_ids)
    dtype = P["bert.pooler.dense.weight"].dtype
    ext = extended_mask(attention_mask, dtype)
    x = embeddings(P, cfg, input_ids, token_type_ids, position_ids, drop)
    if img_feats is not None:
        im = F.linear(img_feats, P["bert.img_embedding.weight"], P["bert.img_embedding.bias"])
        im = im + F.linear(img_location_embeddings, P["bert.location_embeds.weight"],
                           P["bert.location_embeds.bias"])
        if cfg.use_img_layernorm:
            im = bert_layer_norm(im, P["bert.LayerNorm.weight"], P["bert.LayerNorm.bias"],
                                 cfg.img_layer_norm_eps)
        if drop:
            im = drop(im, cfg.hidden_dropout_prob)
        x = torch.cat((x, im), 1)
    hidden = [x]
    for i in range(cfg.num_hidden_layers):
        x = encoder_layer(P, cfg, i, x, ext, drop)
        hidden.append(x)
    pooled = torch.tanh(F.linear(x[:, 0], P["bert.pooler.dense.weight"], P["bert.pooler.dense.bias"]))
    if return_hidden:
        return x, pooled, hidden
    return x, pooled


# --------------------------------------------------------------------------------------------
# pre-training heads
# --------------------------------------------------------------------------------------------
def mlm_logits(P: Params, cfg, h):
    """BertOnlyMLMHead, hf/modeling_bert.py:470-513; decoder weight tied to the word embeddings -- unless
    resize_embeddings has replaced the embedding table (vs/encoder.py:337-345 does not re-tie): the decoder then keeps the
    old table as a parameter of its own, here the optional entry "mlmhead.predictions.decoder.weight"."""
    t = F.linear(h, P["mlmhead.predictions.transform.dense.weight"], P["mlmhead.predictions.transform.dense.bias"])
    t = bert_layer_norm(gelu_erf(t), P["mlmhead.predictions.transform.LayerNorm.weight"],
                        P["mlmhead.predictions.transform.LayerNorm.bias"], cfg.layer_norm_eps)
    dec = P.get("mlmhead.predictions.decoder.weight", P["bert.embeddings.word_embeddings.weight"])
    return F.linear(t, dec) + P["mlmhead.predictions.bias"]


def pretrain_forward(P: Params, cfg: OracleConfig, input_ids, token_type_ids=None, attention_mask=None,
                     labels=None, token_labels=None, position_ids=None, img_feats=None,
                     img_location_embeddings=None, next_action=None, drop=None, return_aux=False):
    """PreTrainOscar.forward, vs/encoder.py:347-441.  Returns the 7-tuple
    (loss, mask_loss, next_loss, token_loss, words_accuracy, action_accuracy, token_accuracy)."""
    seq, pooled = bert_img_forward(P, cfg, input_ids, token_type_ids, attention_mask, position_ids,
                                   img_feats, img_location_embeddings, drop)
    V, Dc = cfg.vocab_size, cfg.detector_classes
    pred_scores = mlm_logits(P, cfg, seq)                                     # :377
    tok_prob = torch.softmax(F.linear(seq, P["token_head.0.weight"], P["token_head.0.bias"]), -1)  # :323-326
    token_loss = F.cross_entropy(tok_prob.view(-1, Dc), token_labels.view(-1), ignore_index=-1)    # :381-385
    mask_loss = F.cross_entropy(pred_scores.view(-1, V), labels.view(-1), ignore_index=-1)         # :387-389
    act = F.log_softmax(F.linear(pooled, P["next_action.linear.weight"], P["next_action.linear.bias"]), -1)  # :152-158
    next_loss = 0                                                                                  # :392 (a Python int without next_action)
    if next_action is not None:
        next_loss = F.cross_entropy(act, next_action, ignore_index=-1)                             # :393-394
    loss = mask_loss + next_loss + token_loss                                                      # :396
    pa = act.argmax(1)
    pw = pred_scores.argmax(2)
    pt = tok_prob.argmax(2)
    pw[labels == -1] = -1                                                                          # :402
    ign_w = (labels == -1).sum()
    words_left = (labels.numel() - ign_w).float()
    words_acc = ((pw == labels).sum() - ign_w) / words_left                                        # :408-410
    action_acc = 0                                                                                 # :417-418
    if next_action is not None:
        action_acc = (pa == next_action).sum().float() / pa.shape[0]                               # :412-416
    pt[token_labels == -1] = -1                                                                    # :420
    ign_t = (token_labels == -1).sum()
    tokens_left = (pt.numel() - ign_t).float()
    token_acc = ((pt == token_labels).sum().float() - ign_t) / tokens_left                         # :428-431
    out = (loss, mask_loss, next_loss, token_loss, words_acc, action_acc, token_acc)
    if return_aux:
        return out, dict(sequence_output=seq, pooled_output=pooled, prediction_scores=pred_scores,
                         token_prob=tok_prob, action_scores=act, predicted_action=pa,
                         predicted_words=pw, predicted_tokens=pt)
    return out


# --------------------------------------------------------------------------------------------
# fine-tuning heads
# --------------------------------------------------------------------------------------------
def soft_dot_attention(w_in, w_out, h, context, mask=None, output_tilde=True, output_prob=True):
    """SoftDotAttention.forward, vs/agent_models.py:328-357 (bias-free linears)."""
    target = F.linear(h, w_in)[:, :, None]
    logit = torch.bmm(context, target).squeeze(2)
    attn = logit
    if mask is not None:
        attn = attn.masked_fill(mask.bool(), -float("inf"))
        logit = attn  # the reference masks in place, so the returned logit is masked too (:343)
    attn = torch.softmax(attn, dim=1)
    weighted = torch.bmm(attn[:, None, :], context).squeeze(1)
    if not output_prob:
        attn = logit
    if output_tilde:
        return torch.tanh(F.linear(torch.cat((weighted, h), 1), w_out)), attn
    return weighted, attn


def candidate_scores_ce(w_in, h_tilde, cand_feat, cand_len, target, ignore_index=-100):
    """Candidate-viewpoint scorer + loss: vs/agent_models.py:426 (candidate_att_layer with
    output_prob=False -> raw bilinear logits) then vs/agent.py:396-408 (masked_fill -inf beyond
    candidate_leng, CrossEntropyLoss(ignore_index=-100, reduction mean), argmax)."""
    logit = torch.bmm(cand_feat, F.linear(h_tilde, w_in)[:, :, None]).squeeze(2)
    C = cand_feat.size(1)
    invalid = torch.arange(C, device=cand_len.device)[None, :] >= cand_len[:, None]
    logit = logit.masked_fill(invalid, -float("inf"))
    loss = F.cross_entropy(logit, target, ignore_index=ignore_index)
    return logit, loss, logit.argmax(1)


FORWARD_ACTION = 4   # tb/agent.py:71-80: model_actions = [left, right, up, down, forward, <end>, <start>, <ignore>]


def turn_based_head(w_in, w_out, w_a, b_a, h, ctx, ctx_mask, no_forward, target, ignore_index=7):
    """tb/agent_models.py:253-274,297,316-318 + tb/agent.py:150-152,316-327: ctx attention ->
    h_tilde -> decoder2action logits [B,6]; logit[:, 4] (model_actions.index("forward"), tb/agent.py:71-80) -> -inf where
    no navigable location; CE(ignore_index=7)."""
    h_tilde, _ = soft_dot_attention(w_in, w_out, h, ctx, ctx_mask)
    logit = F.linear(h_tilde, w_a, b_a)
    if no_forward is not None:
        logit = logit.clone()
        logit[:, FORWARD_ACTION] = logit[:, FORWARD_ACTION].masked_fill(no_forward.bool(), -float("inf"))
    loss = F.cross_entropy(logit, target, ignore_index=ignore_index)
    return h_tilde, logit, loss, logit.argmax(1)


def qa_head(w1, b1, w2, b2, h_tilde, target, valid, pos_weight=5.0):
    """question_linear (vs/agent_models.py:455-459,503) + BCEWithLogitsLoss(pos_weight=5,
    reduction='none') masked mean (vs/classifier/agent.py:158-163,493-503)."""
    q = F.linear(torch.tanh(F.linear(h_tilde, w1, b1)), w2, b2)
    per = F.binary_cross_entropy_with_logits(
        q.squeeze(1), target.float(), reduction="none",
        pos_weight=torch.tensor(pos_weight, dtype=q.dtype, device=q.device))
    v = valid.to(per.dtype)
    return q, (per * v).sum() / v.sum().clamp(min=1.0)


# --------------------------------------------------------------------------------------------
# the agents' recurrent middle and rollouts
# --------------------------------------------------------------------------------------------
def lstm_cell(x, h, c, w_ih, w_hh, b_ih, b_hh):
    """nn.LSTMCell / one step of nn.LSTM (gate order i, f, g, o): vs/agent_models.py:380,414 ; :223-230."""
    g = F.linear(x, w_ih, b_ih) + F.linear(h, w_hh, b_hh)
    i, f, gg, o = g.chunk(4, 1)
    c1 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
    return torch.sigmoid(o) * torch.tanh(c1), c1


def lstm_sequence(x, lengths, w_ih, w_hh, b_ih, b_hh):
    """pack_padded_sequence -> nn.LSTM(batch_first) -> pad_packed_sequence, vs/agent_models.py:283-305, as an explicit loop:
    sample b advances while t < lengths[b]; ctx is 0 behind its end and as long as the longest sample; (h, c) are the last
    valid states.  x [B, T, in] -> ctx [B, max(lengths), H], h [B, H], c [B, H]."""
    B, T, _ = x.shape
    Hh = w_hh.size(1)
    lengths = torch.as_tensor(lengths, device=x.device)
    h = x.new_zeros(B, Hh)
    c = x.new_zeros(B, Hh)
    outs = []
    for t in range(int(lengths.max())):
        h1, c1 = lstm_cell(x[:, t], h, c, w_ih, w_hh, b_ih, b_hh)
        live = (t < lengths)[:, None]
        h, c = torch.where(live, h1, h), torch.where(live, c1, c)
        outs.append(torch.where(live, h1, torch.zeros_like(h1)))
    return torch.stack(outs, 1), h, c


def oscar_encoder_forward(P: Params, cfg, A: Params, inputs, lengths, mask, token_type_ids=None, position_ids=None, drop=None,
                          agent_drop=None, p_agent=0.5):
    """OscarEncoder.forward, vs/agent_models.py:254-310 (hidden == decoder hidden: c_t is passed through, :300-301)."""
    att_mask = ~mask                                                                               # :268 (uint8 -> 255, F6)
    seq, _ = bert_img_forward(P, cfg, inputs, token_type_ids, att_mask, position_ids, drop=drop)   # :270-277
    ctx, h_t, c_t = lstm_sequence(seq, lengths, A["lstm.weight_ih_l0"], A["lstm.weight_hh_l0"], A["lstm.bias_ih_l0"], A["lstm.bias_hh_l0"])
    decoder_init = torch.tanh(F.linear(h_t, A["encoder_lstm2decoder_ht.weight"], A["encoder_lstm2decoder_ht.bias"]))   # :299
    if agent_drop is not None:
        ctx = agent_drop(ctx, p_agent)                                                             # :307
    return ctx, decoder_init, c_t


def attn_decoder_step_vs(D: Params, action, feature, cand_feat, h_0, prev_h1, c_0, ctx, ctx_mask=None, drop=None, p=0.5):
    """AttnDecoderLSTM.forward of viewpoint select, vs/agent_models.py:384-428 -> (h_1, c_1, logit, h_tilde)."""
    dr = (lambda t: drop(t, p)) if drop is not None else (lambda t: t)
    action_embeds = dr(torch.tanh(F.linear(action, D["embedding.0.weight"], D["embedding.0.bias"])))
    attn_feat, _ = soft_dot_attention(D["feat_att_layer.linear_in.weight"], None, dr(prev_h1), feature, output_tilde=False)
    h_1, c_1 = lstm_cell(torch.cat((action_embeds, attn_feat), 1), prev_h1, c_0, D["lstm.weight_ih"], D["lstm.weight_hh"],
                         D["lstm.bias_ih"], D["lstm.bias_hh"])
    h_tilde, _ = soft_dot_attention(D["attention_layer.linear_in.weight"], D["attention_layer.linear_out.weight"], dr(h_1), ctx, ctx_mask)
    _, logit = soft_dot_attention(D["candidate_att_layer.linear_in.weight"], None, dr(h_tilde), cand_feat, output_tilde=False,
                                  output_prob=False)
    return h_1, c_1, logit, h_tilde


def attn_decoder_step_tb(D: Params, action, feature, h_0, c_0, ctx, ctx_mask=None, drop=None, p=0.5):
    """AttnDecoderLSTM.forward of turn based, tb/agent_models.py:299-319 -> (h_1, c_1, alpha, logit)."""
    dr = (lambda t: drop(t, p)) if drop is not None else (lambda t: t)
    x = dr(torch.cat((D["embedding.weight"][action.view(-1)], feature), 1))
    h_1, c_1 = lstm_cell(x, h_0, c_0, D["lstm.weight_ih"], D["lstm.weight_hh"], D["lstm.bias_ih"], D["lstm.bias_hh"])
    h_tilde, alpha = soft_dot_attention(D["attention_layer.linear_in.weight"], D["attention_layer.linear_out.weight"], dr(h_1), ctx, ctx_mask)
    return h_1, c_1, alpha, F.linear(h_tilde, D["decoder2action.weight"], D["decoder2action.bias"])


def _feedback(logit, target, feedback):
    if feedback == "teacher":
        return target
    if feedback == "argmax":
        return logit.max(1)[1].detach()
    raise ValueError("the oracle restates the deterministic feedback modes (teacher, argmax)")


def rollout_vs(P, cfg, A, D, batch, trace, episode_len, feedback="teacher", ignore_index=-100):
    """Agent.rollout of viewpoint select, vs/agent.py:334-474, over a RECORDED trace (observations per step instead of the
    simulator): loss = sum_t CE(masked logit_t, target_t) / episode_len; a finished episode's targets are ignored."""
    ctx, h_t, c_t = oscar_encoder_forward(P, cfg, A, batch["input_ids"], batch["lengths"], batch["mask"], batch["token_type_ids"])
    ctx_mask = batch["mask"][:, :ctx.size(1)]
    h1 = h_t
    ended = torch.zeros(ctx.size(0), dtype=torch.bool, device=ctx.device)
    loss, logits = 0.0, []
    for t in range(episode_len):
        h_t, c_t, logit, h1 = attn_decoder_step_vs(D, trace["input_a_t"][t], trace["f_t"][t], trace["cand_feat"][t], h_t, h1, c_t, ctx, ctx_mask)
        leng = trace["cand_len"][t]
        invalid = torch.arange(logit.size(1), device=leng.device)[None, :] >= leng[:, None]       # utils.length2mask, :396-403
        logit = logit.masked_fill(invalid, -float("inf"))
        target = torch.where(ended, torch.full_like(trace["target"][t], ignore_index), trace["target"][t])   # _teacher_action(obs, ended)
        if bool((target != ignore_index).any()):    # (a recorded trace can ignore every sample of a step; the live teacher cannot)
            loss = loss + F.cross_entropy(logit, target, ignore_index=ignore_index)                # :405-409
        a_t = _feedback(logit, target, feedback)                                                   # :411-423
        ended = ended | (a_t == leng - 1) | (a_t == ignore_index)                                  # :429-441
        logits.append(logit)
        if bool(ended.all()):                                                                      # :466-468
            break
    return loss / episode_len, logits                                                              # :472-474


def rollout_tb(P, cfg, A, D, batch, trace, episode_len, feedback="teacher", ignore_index=7, start_action=6, end_action=5):
    """Agent.rollout of turn based, tb/agent.py:259-392, over a recorded trace."""
    ctx, h_t, c_t = oscar_encoder_forward(P, cfg, A, batch["input_ids"], batch["lengths"], batch["mask"], batch["token_type_ids"])
    ctx_mask = batch["mask"][:, :ctx.size(1)]
    B = ctx.size(0)
    a_t = torch.full((B,), start_action, dtype=torch.long, device=ctx.device)                      # :287-290
    ended = torch.zeros(B, dtype=torch.bool, device=ctx.device)
    loss, logits = 0.0, []
    for t in range(episode_len):
        h_t, c_t, _, logit = attn_decoder_step_tb(D, a_t.view(-1, 1), trace["f_t"][t], h_t, c_t, ctx, ctx_mask)
        logit = logit.clone()
        logit[:, FORWARD_ACTION] = logit[:, FORWARD_ACTION].masked_fill(trace["no_forward"][t].bool(), -float("inf"))   # :316-318
        target = torch.where(ended, torch.full_like(trace["target"][t], ignore_index), trace["target"][t])
        if bool((target != ignore_index).any()):
            loss = loss + F.cross_entropy(logit, target, ignore_index=ignore_index)
        a_t = _feedback(logit, target, feedback)
        ended = ended | (a_t == end_action)                                                        # :343-346
        logits.append(logit)
        if bool(ended.all()):                                                                      # :388-390
            break
    return loss / episode_len, logits


def init_agent_params(seed: int, hidden: int = 512, feat: int = 2052, angle: int = 4, emb: int = 64, tb_feat: int = 2048, tb_emb: int = 32,
                      scale: float = 0.05):
    """Seeded weights of OscarEncoder's LSTM part (A), the viewpoint-select decoder (Dv) and the turn-based decoder (Dt); key names =
    the reference modules' state_dict keys."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: scale * torch.randn(*s, generator=g)  # noqa: E731
    A = {"lstm.weight_ih_l0": r(4 * hidden, 768), "lstm.weight_hh_l0": r(4 * hidden, hidden), "lstm.bias_ih_l0": r(4 * hidden),
         "lstm.bias_hh_l0": r(4 * hidden), "encoder_lstm2decoder_ht.weight": r(hidden, hidden), "encoder_lstm2decoder_ht.bias": r(hidden),
         "encoder_lstm2decoder_ct.weight": r(hidden, hidden), "encoder_lstm2decoder_ct.bias": r(hidden)}
    Dv = {"embedding.0.weight": r(emb, angle), "embedding.0.bias": r(emb), "lstm.weight_ih": r(4 * hidden, emb + feat),
          "lstm.weight_hh": r(4 * hidden, hidden), "lstm.bias_ih": r(4 * hidden), "lstm.bias_hh": r(4 * hidden),
          "feat_att_layer.linear_in.weight": r(feat, hidden), "feat_att_layer.linear_out.weight": r(hidden, hidden + feat),
          "attention_layer.linear_in.weight": r(hidden, hidden), "attention_layer.linear_out.weight": r(hidden, 2 * hidden),
          "candidate_att_layer.linear_in.weight": r(feat, hidden), "candidate_att_layer.linear_out.weight": r(hidden, hidden + feat)}
    Dt = {"embedding.weight": r(8, tb_emb), "lstm.weight_ih": r(4 * hidden, tb_emb + tb_feat), "lstm.weight_hh": r(4 * hidden, hidden),
          "lstm.bias_ih": r(4 * hidden), "lstm.bias_hh": r(4 * hidden), "attention_layer.linear_in.weight": r(hidden, hidden),
          "attention_layer.linear_out.weight": r(hidden, 2 * hidden), "decoder2action.weight": r(6, hidden), "decoder2action.bias": r(6)}
    return A, Dv, Dt


def init_head_params(seed: int, hidden: int = 512, feat: int = 2052) -> Params:
    """Seeded weights of the fine-tuning heads (shapes: vs/agent_models.py:382,455-459;
    tb/agent_models.py:243-246,297)."""
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: 0.05 * torch.randn(*s, generator=g)  # noqa: E731
    return {
        "candidate_att_layer.linear_in.weight": r(feat, hidden),
        "attention_layer.linear_in.weight": r(hidden, hidden),
        "attention_layer.linear_out.weight": r(hidden, 2 * hidden),
        "decoder2action.weight": r(6, hidden),
        "decoder2action.bias": r(6),
        "question_linear.0.weight": r(hidden // 2, hidden),
        "question_linear.0.bias": r(hidden // 2),
        "question_linear.2.weight": r(1, hidden // 2),
        "question_linear.2.bias": r(1),
    }


# --------------------------------------------------------------------------------------------
# the step after backward
# --------------------------------------------------------------------------------------------
def adamw_update(p, g, m, v, step: int, lr: float, beta1=0.9, beta2=0.999, eps=1e-6, weight_decay=0.0, correct_bias=True):
    """One AdamW update of one tensor, transformers/pytorch_transformers/optimization.py:165-187: moments, eps OUTSIDE
    the square root, optional bias correction folded into the step size, then the decoupled decay with the
    un-corrected lr applied to the already updated weights.  ``step`` is the 1-based update count.
    Returns new (p, m, v); inputs are not modified."""
    m = m * beta1 + (1.0 - beta1) * g
    v = v * beta2 + (1.0 - beta2) * g * g
    denom = v.sqrt() + eps
    step_size = lr
    if correct_bias:
        step_size = step_size * math.sqrt(1.0 - beta2 ** step) / (1.0 - beta1 ** step)
    p = p - step_size * (m / denom)
    if weight_decay > 0.0:
        p = p - lr * weight_decay * p
    return p, m, v


def warmup_linear(step: int, warmup_steps: int, t_total: int) -> float:
    """WarmupLinearSchedule.lr_lambda, optimization.py:59-62."""
    if step < warmup_steps:
        return float(step) / float(max(1, warmup_steps))
    return max(0.0, float(t_total - step) / float(max(1.0, t_total - warmup_steps)))


def clip_grad_norm(grads, max_norm: float):
    """torch.nn.utils.clip_grad_norm_ (L2) as called at vs/agent.py:514-518: returns (total norm, scaled grads)."""
    total = torch.sqrt(sum((g.double() ** 2).sum() for g in grads)).float()
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    return total, [g * coef for g in grads]


# --------------------------------------------------------------------------------------------
# the step before forward
# --------------------------------------------------------------------------------------------
def mlm_mask(inputs, token_classes, special_mask, chosen, as_mask, as_random, random_words, mask_id, pad_id):
    """PretrainDataset._mask_tokens, vs/data_loader_pretrain.py:549-612, as a function of its random draws:
    ``chosen`` = bernoulli(0.15), ``as_mask`` = bernoulli(0.8), ``as_random`` = bernoulli(0.5) (bool tensors),
    ``random_words`` = randint(len(tokenizer)); ``token_classes`` None = masked_token_prediction off.
    Returns (inputs', labels, attention_mask)."""
    inputs = inputs.clone()
    labels = inputs.clone()
    masked = chosen & ~special_mask                                   # :555-566 (probability 0 on special tokens)
    tc = None if token_classes is None else token_classes != -1
    if tc is not None:
        masked = masked | tc                                          # :568-571
    attention = inputs != pad_id                                      # :573-575
    labels[~masked] = -1                                              # :577
    if tc is not None:
        labels[tc] = -1                                               # :579-580
    replaced = as_mask & masked                                       # :583-586
    inputs[replaced] = mask_id                                        # :588-590
    if tc is not None:
        replaced = replaced | tc                                      # :591-597
        inputs[tc] = mask_id
    rnd = as_random & masked & ~replaced                              # :600-604
    inputs[rnd] = random_words[rnd]                                   # :606-610
    return inputs, labels, attention


def preprocess_item(masked, token_classes, img_features, location_embeddings, next_action, max_img_seq_length,
                    masked_token_prediction=True, no_action_grounding=False):
    """PretrainDataset._preprocess_item, vs/data_loader_pretrain.py:636-712, after its _mask_tokens call (``masked`` = that call's
    (inputs, labels, attention_mask)) and its feature read (``img_features`` [n, D], ``location_embeddings`` [n, Dl])."""
    inputs, labels, attention = masked
    R = max_img_seq_length
    n = img_features.shape[0]
    attention = [int(a) for a in attention.tolist()]
    if n > R:                                                          # :656-665 keep the LAST R rows
        img_features, location_embeddings = img_features[n - R:] if R > 0 else img_features[n:], \
            location_embeddings[n - R:] if R > 0 else location_embeddings[n:]
        if R > 0:
            attention = attention + [1] * img_features.shape[0]
    else:                                                              # :666-690 zero padding, attention 0 on it
        if R > 0:
            attention = attention + [1] * n
        img_features = torch.cat([img_features, torch.zeros(R - n, img_features.shape[1], dtype=img_features.dtype)], 0)
        location_embeddings = torch.cat([location_embeddings, torch.zeros(R - n, location_embeddings.shape[1],
                                                                          dtype=location_embeddings.dtype)], 0)
        if R > 0:
            attention = attention + [0] * (R - n)
    tail = [-1] * img_features.shape[0]
    return dict(input_ids=inputs, labels=torch.tensor(labels.tolist() + tail, dtype=torch.long),                       # :692-693
                token_labels=torch.tensor(token_classes.tolist() + tail, dtype=torch.long) if masked_token_prediction else None,   # :697-702
                attention_mask=attention, img_feats=img_features, img_location_embeddings=location_embeddings,
                next_action=-1 if no_action_grounding else next_action)                                                 # :695-696


# --------------------------------------------------------------------------------------------
# helpers for tests / bench
# --------------------------------------------------------------------------------------------
def clone_params(P: Params, requires_grad=True, dtype=None, device=None) -> Params:
    out = {}
    for k, v in P.items():
        t = v.detach().clone()
        if dtype is not None:
            t = t.to(dtype)
        if device is not None:
            t = t.to(device)
        out[k] = t.requires_grad_(requires_grad)
    return out


def train_dropout(x, p):
    return F.dropout(x, p, training=True) if p > 0 else x


def pretrain_step(P: Params, cfg: OracleConfig, batch: dict, drop=None):
    """One forward+backward of the pre-training path (the unit BASELINE.json's metric counts).
    Returns (7-tuple of floats, {name: grad})."""
    for v in P.values():
        v.grad = None
    out = pretrain_forward(P, cfg, drop=drop, **batch)
    out[0].backward()
    return tuple(float(o.detach()) for o in out), {k: v.grad for k, v in P.items()}
