"""Device-resident rollouts of the reference's agents.

The reference's fine-tuning hot loop (vs/agent.py:334-474, tb/agent.py:259-392) rebuilds every decoder step's inputs in
numpy from simulator observations, and reads the chosen actions back with ``a_t.cpu()`` / ``.item()`` -- a host <-> device
sync per step, ten to forty per iteration.  Here an episode runs over a TRACE held on the device (what the simulator
would have delivered at each step of the recorded path: view features, candidate features and counts, heading features,
teacher actions), the action feedback (teacher forcing / argmax / sampling with the counter RNG) and the ``ended``
bookkeeping are one small kernel per step (``vt_rollout_feedback``), and nothing in the episode touches the host: the
whole iteration -- encoder, 511-step LSTM, every decoder step, loss, backward -- is a fixed launch sequence that
``visitron_b200.graph.GraphedStep`` captures in one CUDA graph.

A trace is exact for teacher forcing (the recorded path IS the teacher's path); under argmax / sample feedback the chosen
action feeds the decoder and the ``ended`` state, while the observations stay the recorded ones (there is no simulator here;
the Matterport3D simulator is out of scope).  Unlike the reference the loop never exits early when every
episode has ended (`if ended.all(): break` is a sync): finished episodes carry the ignore target and a step whose samples are all
finished adds exactly 0 to the loss (and no gradient), so the result is the reference's.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from . import lib, modeling
from .agent_heads import masked_cross_entropy
from .lib import check, init, ptr, stream_ptr

FEEDBACK = {"teacher": 0, "argmax": 1, "sample": 2}


def _feedback(masked_logit, target, stop_len, stop_const, ignore_index, mode, seed, step, ended, next_target):
    """One launch: (a_t, effective target of the next step); `ended` is updated in place.  No gradient flows through actions."""
    B, C = masked_logit.shape
    a = torch.empty(B, device=masked_logit.device, dtype=torch.int64)
    nxt = None if next_target is None else torch.empty_like(next_target)
    check(init().vt_rollout_feedback(ptr(masked_logit), masked_logit.stride(0), B, C, ptr(target), ptr(stop_len), stop_const, ignore_index, mode,
                                     seed, 0x7011, step, ptr(ended), ptr(a), ptr(next_target), ptr(nxt), stream_ptr()), "vt_rollout_feedback")
    return a, nxt


def _check_trace(trace: Dict[str, torch.Tensor], keys, episode_len: int):
    for k in keys:
        if k not in trace:
            raise lib.VtError("rollout trace lacks %r" % k)
        if not trace[k].is_cuda or trace[k].size(0) < episode_len:
            raise lib.VtError("trace[%r] must be a device tensor with at least episode_len = %d steps in dim 0" % (k, episode_len))


def viewpoint_rollout(encoder, decoder, batch: Dict[str, torch.Tensor], trace: Dict[str, torch.Tensor], episode_len: int,
                      feedback: str = "teacher", ignore_index: int = -100) -> Tuple[torch.Tensor, List[torch.Tensor]]:
    """Agent.rollout of viewpoint select (vs/agent.py:334-474) -> (loss, masked logits per step).

    batch: input_ids [b,511], token_type_ids, mask (uint8 padding mask as _sort_batch builds it), lengths (list or tensor).
    trace: input_a_t [T,b,4], f_t [T,b,36,2052], cand_feat [T,b,C,2052], cand_len [T,b] int64, target [T,b] int64."""
    _check_trace(trace, ("input_a_t", "f_t", "cand_feat", "cand_len", "target"), episode_len)
    mode = FEEDBACK[feedback]
    ctx, h_t, c_t = encoder(batch["input_ids"], batch["lengths"], batch["mask"], token_type_ids=batch.get("token_type_ids"))
    ctx_mask = batch["mask"][:, :ctx.size(1)]
    h1 = h_t
    B = ctx.size(0)
    ended = torch.zeros(B, device=ctx.device, dtype=torch.uint8)
    seed = modeling._draw_seed() if mode == 2 else 0
    target = trace["target"][0]
    loss, logits = None, []
    for t in range(episode_len):
        h_t, c_t, logit, h1 = decoder(trace["input_a_t"][t], trace["f_t"][t], trace["cand_feat"][t], h_t, h1, c_t, ctx, ctx_mask)
        l, masked, _ = masked_cross_entropy(logit, target, valid_len=trace["cand_len"][t], ignore_index=ignore_index, empty_is_zero=True)
        loss = l if loss is None else loss + l
        logits.append(masked)
        nxt = trace["target"][t + 1] if t + 1 < episode_len else None
        _, target = _feedback(masked, target, trace["cand_len"][t], 0, ignore_index, mode, seed, t, ended, nxt)
    return loss / episode_len, logits


def turn_based_rollout(encoder, decoder, batch: Dict[str, torch.Tensor], trace: Dict[str, torch.Tensor], episode_len: int,
                       feedback: str = "teacher", ignore_index: int = 7, start_action: int = 6, end_action: int = 5
                       ) -> Tuple[torch.Tensor, List[torch.Tensor]]:
    """Agent.rollout of turn based (tb/agent.py:259-392) -> (loss, masked logits per step).

    trace: f_t [T,b,2048], no_forward [T,b] bool (no navigable location: "forward" is impossible), target [T,b] int64."""
    _check_trace(trace, ("f_t", "no_forward", "target"), episode_len)
    mode = FEEDBACK[feedback]
    ctx, h_t, c_t = encoder(batch["input_ids"], batch["lengths"], batch["mask"], token_type_ids=batch.get("token_type_ids"))
    ctx_mask = batch["mask"][:, :ctx.size(1)]
    B = ctx.size(0)
    a_t = torch.full((B,), start_action, device=ctx.device, dtype=torch.int64)
    ended = torch.zeros(B, device=ctx.device, dtype=torch.uint8)
    seed = modeling._draw_seed() if mode == 2 else 0
    target = trace["target"][0]
    loss, logits = None, []
    for t in range(episode_len):
        h_t, c_t, _, logit = decoder(a_t.view(-1, 1), trace["f_t"][t], h_t, c_t, ctx, ctx_mask)
        l, masked, _ = masked_cross_entropy(logit, target, forbid=trace["no_forward"][t], ignore_index=ignore_index, empty_is_zero=True)
        loss = l if loss is None else loss + l
        logits.append(masked)
        nxt = trace["target"][t + 1] if t + 1 < episode_len else None
        a_t, target = _feedback(masked, target, None, end_action, ignore_index, mode, seed, t, ended, nxt)
    return loss / episode_len, logits
