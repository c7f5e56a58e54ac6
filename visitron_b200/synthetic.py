"""Synthetic batches with the reference's exact tensor shapes / dtypes / value ranges.

The Matterport3D simulator and the precomputed Faster R-CNN features are not available
(BASELINE.json), so every workload is fed from these generators.  Shapes and value ranges follow
the reference data loaders:

* pre-training batch  -- keys of ``tasks/viewpoint_select/data_loader_pretrain.py:703-711``:
  ``input_ids, labels, token_labels, attention_mask, img_feats, img_location_embeddings,
  next_action`` (no ``token_type_ids``), text padded to 511 (``:89-92,210``), 36 views x 5 boxes
  = 180 real regions padded to ``max_img_seq_length`` 256 (``:615-634,659-690``), BERT 80/10/10
  masking (``:555-596``), features = ``torch.rand`` like the reference's ``--debug`` stand-in
  (``:620-621``), location embeddings as ``build_viewpoint_loc_embedding`` (``:25-49``).
* fine-tuning encoder batch -- ``tasks/viewpoint_select/agent.py:159-181``: ids with no zeros
  (pads are [UNK]=100), ``segment_ids`` in {0,1}, ``mask`` uint8 all-zero and the
  encoder is called with ``attention_mask = ~mask`` (== 255).
"""
from __future__ import annotations

import math
from typing import Dict

import torch

CLS, SEP, MASK, UNK = 101, 102, 103, 100


def location_embedding_table() -> torch.Tensor:
    """[36 (current view), 36 (absolute view), 128] heading/elevation sin-cos table
    (tasks/viewpoint_select/data_loader_pretrain.py:25-49)."""
    inc = math.pi / 6.0
    tab = torch.zeros(36, 36, 128)
    for view in range(36):
        for a in range(36):
            rel = (a - view) % 12 + (a // 12) * 12
            h, e = (rel % 12) * inc, (rel // 12 - 1) * inc
            tab[view, a, 0:32] = math.sin(h)
            tab[view, a, 32:64] = math.cos(h)
            tab[view, a, 64:96] = math.sin(e)
            tab[view, a, 96:128] = math.cos(e)
    return tab


def pretrain_batch(batch: int, seed: int = 88, text_len: int = 511, regions: int = 256,
                   img_dim: int = 2054, vocab: int = 30522, detector_classes: int = 1601,
                   action_space: int = 36, full_text: bool = False) -> Dict[str, torch.Tensor]:
    """NDH-shaped pre-training batch on the CPU (callers move / pin it)."""
    g = torch.Generator().manual_seed(seed)
    T, R, S = text_len, regions, text_len + regions
    lo_id = min(1000, vocab // 2)
    ids = torch.zeros(batch, T, dtype=torch.long)
    labels = torch.full((batch, S), -1, dtype=torch.long)
    token_labels = torch.full((batch, S), -1, dtype=torch.long)
    mask = torch.zeros(batch, S, dtype=torch.long)
    n_real = (regions * 45) // 64  # 180 of 256
    views = max(1, n_real // 5)
    tab = location_embedding_table()
    img = torch.zeros(batch, R, img_dim)
    loc = torch.zeros(batch, R, 128)
    for b in range(batch):
        lt = T if full_text else int(torch.randint(max(8, (2 * T) // 5), T + 1, (1,), generator=g))
        row = torch.randint(lo_id, vocab, (lt,), generator=g)
        row[0] = CLS
        row[lt - 1] = SEP
        # region-label tokens: the tail of the text, ~ 10-23 % of it, detector class targets
        n_tok = int(torch.randint(max(1, lt // 10), max(2, lt // 4), (1,), generator=g))
        tok_pos = torch.arange(lt - 1 - n_tok, lt - 1)
        text_pos = torch.arange(1, lt - 1 - n_tok)
        # BERT masking on the dialog part: 15 % of positions; 80 % [MASK], 10 % random, 10 % kept
        pick = text_pos[torch.rand(text_pos.numel(), generator=g) < 0.15]
        if pick.numel() == 0:
            pick = text_pos[:1]
        labels[b, pick] = row[pick]
        r = torch.rand(pick.numel(), generator=g)
        rnd = torch.randint(lo_id, vocab, (pick.numel(),), generator=g)
        row[pick] = torch.where(r < 0.8, torch.full_like(rnd, MASK), torch.where(r < 0.9, rnd, row[pick]))
        # masked-token prediction: ~ half of the region-label tokens are forced to [MASK]
        tpick = tok_pos[torch.rand(tok_pos.numel(), generator=g) < 0.5]
        if tpick.numel() == 0:
            tpick = tok_pos[:1]
        token_labels[b, tpick] = torch.randint(0, detector_classes, (tpick.numel(),), generator=g)
        row[tpick] = MASK
        ids[b, :lt] = row
        mask[b, :lt] = 1
        mask[b, T:T + n_real] = 1
        img[b, :n_real] = torch.rand(n_real, img_dim, generator=g)
        cur = int(torch.randint(0, 36, (1,), generator=g))
        view_of_region = (torch.arange(n_real) // max(1, n_real // views)).clamp(max=35)
        loc[b, :n_real] = tab[cur, view_of_region]
    next_action = torch.randint(0, action_space, (batch,), generator=g)
    return dict(input_ids=ids, labels=labels, token_labels=token_labels, attention_mask=mask,
                img_feats=img, img_location_embeddings=loc, next_action=next_action)


def finetune_encoder_batch(batch: int, seed: int = 88, text_len: int = 511, vocab: int = 30522,
                           valid_len: int | None = None) -> Dict[str, torch.Tensor]:
    """What ``OscarEncoder.forward`` hands to the encoder (vs/agent_models.py:267-275)."""
    g = torch.Generator().manual_seed(seed)
    lo_id = min(1000, vocab // 2)
    ids = torch.randint(lo_id, vocab, (batch, text_len), generator=g)
    ids[:, 0] = CLS
    for b in range(batch):
        n = valid_len if valid_len is not None else int(
            torch.randint(max(4, text_len // 4), text_len + 1, (1,), generator=g))
        ids[b, n:] = UNK  # int 0 in a token list -> [UNK]; never pad_token_id 0
    seg = torch.zeros(batch, text_len, dtype=torch.long)
    for b in range(batch):
        cut = int(torch.randint(1, text_len, (1,), generator=g))
        seg[b, cut:] = 1
    mask = torch.zeros(batch, text_len, dtype=torch.uint8)  # all lengths == 511 -> nothing masked
    return dict(input_ids=ids, token_type_ids=seg, attention_mask=~mask)


def viewpoint_step_batch(batch: int, seed: int = 88, candidates: int = 12, feat: int = 2052,
                         hidden: int = 512) -> Dict[str, torch.Tensor]:
    """One decoder-step worth of candidate-scorer inputs (vs/agent.py:186-228,380-408)."""
    g = torch.Generator().manual_seed(seed)
    leng = torch.randint(2, candidates + 1, (batch,), generator=g)
    cand = 0.3 * torch.randn(batch, candidates, feat, generator=g)
    cand = cand * (torch.arange(candidates)[None, :, None] < leng[:, None, None])
    target = (torch.rand(batch, generator=g) * leng).long()
    target[0] = -100  # an already-ended episode: ignore_index (vs/agent.py:128)
    h = torch.randn(batch, hidden, generator=g)
    return dict(h_tilde=h, cand_feat=cand, cand_len=leng, target=target)


def turn_step_batch(batch: int, seed: int = 88, ctx_len: int = 511, hidden: int = 512) -> Dict[str, torch.Tensor]:
    """One turn-based decoder step's head inputs (tb/agent.py:316-327)."""
    g = torch.Generator().manual_seed(seed)
    h = torch.randn(batch, hidden, generator=g)
    ctx = torch.randn(batch, ctx_len, hidden, generator=g)
    lens = torch.randint(ctx_len // 2, ctx_len + 1, (batch,), generator=g)
    ctx_mask = torch.arange(ctx_len)[None, :] >= lens[:, None]
    no_forward = torch.rand(batch, generator=g) < 0.25
    target = torch.randint(0, 6, (batch,), generator=g)
    target = torch.where(no_forward & (target == 4), torch.ones_like(target), target)   # the teacher never goes "forward" (4) where it is impossible
    target[-1] = 7  # <ignore>
    return dict(h=h, ctx=ctx, ctx_mask=ctx_mask, no_forward=no_forward, target=target)


def agent_batch(batch: int, seed: int = 88, text_len: int = 511, vocab: int = 30522, ragged: bool = False) -> Dict[str, torch.Tensor]:
    """What Agent._sort_batch hands to OscarEncoder (vs/agent.py:143-184): ids, segment ids, the uint8 padding mask and the lengths.
    Default = the scripts' situation (no id equals pad_token_id, every length is 511, the mask is all zero);
    ragged=True draws shorter samples with a real padding mask, sorted by descending length as pack_padded_sequence needs."""
    enc = finetune_encoder_batch(batch, seed, text_len=text_len, vocab=vocab)
    g = torch.Generator().manual_seed(seed + 7)
    if ragged:
        lengths = torch.randint(max(2, text_len // 3), text_len + 1, (batch,), generator=g).sort(descending=True).values
        lengths[0] = text_len
    else:
        lengths = torch.full((batch,), text_len, dtype=torch.long)
    mask = (torch.arange(text_len)[None, :] >= lengths[:, None]).to(torch.uint8)
    return dict(input_ids=enc["input_ids"], token_type_ids=enc["token_type_ids"], mask=mask, lengths=lengths)


def viewpoint_trace(steps: int, batch: int, seed: int = 88, candidates: int = 12, feat: int = 2052, views: int = 36,
                    angle: int = 4) -> Dict[str, torch.Tensor]:
    """A recorded viewpoint-select episode (vs/agent.py:186-228,380-441): per step the heading feature, the 36 view features, the
    candidate features (zero beyond candidate_leng, the last valid one = <end> = zeros) and the teacher's action; an episode ends
    when the teacher picks <end>, after which its targets are the ignore id -100."""
    g = torch.Generator().manual_seed(seed)
    leng = torch.randint(2, candidates + 1, (steps, batch), generator=g)
    cand = 0.3 * torch.randn(steps, batch, candidates, feat, generator=g)
    c_idx = torch.arange(candidates)[None, None, :, None]
    cand = cand * (c_idx < (leng - 1)[:, :, None, None])          # the <end> candidate and the padding are zero rows
    f_t = 0.3 * torch.randn(steps, batch, views, feat, generator=g)
    heading = torch.rand(steps, batch, generator=g) * 6.283
    elev = (torch.rand(steps, batch, generator=g) - 0.5)
    a_in = torch.stack([heading.sin(), heading.cos(), elev.sin(), elev.cos()], -1)[..., :angle]
    target = (torch.rand(steps, batch, generator=g) * leng).long().clamp(max=candidates - 1)
    end_at = torch.randint(max(1, steps // 2), steps + 1, (batch,), generator=g)       # the step at which the teacher stops
    for b in range(batch):
        e = int(end_at[b])
        for t in range(steps):
            if t == e - 1:
                target[t, b] = leng[t, b] - 1
            elif t >= e:
                target[t, b] = -100
            elif target[t, b] == leng[t, b] - 1:
                target[t, b] = 0
    return dict(input_a_t=a_in.contiguous(), f_t=f_t, cand_feat=cand, cand_len=leng, target=target)


def turn_trace(steps: int, batch: int, seed: int = 88, feat: int = 2048) -> Dict[str, torch.Tensor]:
    """A recorded turn-based episode (tb/agent.py:301-346): per step the single-view feature, whether "forward" is impossible and the
    teacher's action in [left, right, up, down, forward, <end>]; after <end> the targets are <ignore> = 7."""
    g = torch.Generator().manual_seed(seed)
    f_t = 0.3 * torch.randn(steps, batch, feat, generator=g)
    no_forward = torch.rand(steps, batch, generator=g) < 0.25
    target = torch.randint(0, 5, (steps, batch), generator=g)
    target = torch.where(no_forward & (target == 4), torch.ones_like(target), target)
    end_at = torch.randint(max(1, steps // 2), steps + 1, (batch,), generator=g)
    for b in range(batch):
        e = int(end_at[b])
        target[e - 1, b] = 5
        target[e:, b] = 7
    return dict(f_t=f_t, no_forward=no_forward, target=target)
