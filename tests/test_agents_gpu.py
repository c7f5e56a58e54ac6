"""SURVEY 8f-N1 / N2 on the B200: the cluster LSTM, the decoder steps and the device-resident rollouts against (a) the golden
fixture produced by the unmodified OscarEncoder / AttnDecoderLSTM and (b) the pinned oracle at the path's real sizes
(T = 511, b = 8 / 32), forward and every gradient.  Everything goes through the C ABI."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from visitron_b200 import agent_models as am, graph as vgraph, lib, modeling, rollout, synthetic
    from visitron_b200.agent_models import _LSTMSeqFn

from oracle import visitron_oracle as vo
from parity_common import GRAD_COS, GRAD_REL_L2, LOGIT_TOL, LOSS_RTOL, digest_close, grad_close, load_fixture, make_config

DEV = "cuda"
BF = torch.bfloat16


def _dev(d):
    return {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in d.items()}


def _logit_tol(r):
    """Action logits are bilinear scores cand . (W h~) of magnitude ~10 here; their bf16-level error scales with the ROW's scale,
    so the tolerance is the absolute logit tolerance plus 2 % of the row's largest finite magnitude."""
    fin = torch.isfinite(r)
    return 2 * LOGIT_TOL + 2e-2 * torch.where(fin, r.abs(), torch.zeros_like(r)).max(1, keepdim=True).values


def _logits_close(mine, r, t):
    fin = torch.isfinite(r)
    assert torch.equal(torch.isfinite(mine), fin), t
    err = torch.where(fin, (mine - r).abs(), torch.zeros_like(r))
    assert bool((err <= _logit_tol(r)).all()), (t, float(err.max()))


def _build(oc, P, A, D, space, dropout=0.0):
    bert = modeling.BertImgModelwithLocationEmbeds(make_config(oc))
    bert.load_state_dict({k[len("bert."):]: v.clone() for k, v in P.items()}, strict=True)
    enc = am.OscarEncoder(None, bert, 512, 512, dropout)
    dec = am.AttnDecoderLSTM(4, 64, 512, dropout, 2052) if space == "vs" else am.TurnAttnDecoderLSTM(8, 6, 32, 512, dropout, 2048)
    for mod, W in ((enc, A), (dec, D)):
        sd = mod.state_dict()
        for k, v in W.items():
            assert k in sd and tuple(sd[k].shape) == tuple(v.shape), k
            sd[k] = v.clone()
        mod.load_state_dict(sd, strict=True)
    return enc.to(DEV).train(), dec.to(DEV).train()


@pytest.mark.parametrize("B,T,ragged", [(3, 48, True), (8, 511, False), (9, 130, True), (1, 1, False), (32, 511, True)])
def test_lstm_sequence_kernel_vs_oracle(B, T, ragged):
    """The cluster kernel alone: fp32 x -> bf16 operand -> (ctx, h, c) and all five gradients against the explicit-loop oracle."""
    torch.manual_seed(B * 1000 + T)
    A, _, _ = vo.init_agent_params(3)
    w = [A["lstm." + k].to(DEV).requires_grad_(True) for k in ("weight_ih_l0", "weight_hh_l0", "bias_ih_l0", "bias_hh_l0")]
    x = torch.randn(B, T, 768, device=DEV).to(BF)
    lengths = torch.randint(max(1, T // 3), T + 1, (B,)).sort(descending=True).values if ragged else torch.full((B,), T)
    lengths[0] = T
    xr = x.float().requires_grad_(True)
    wr = [t.detach().clone().requires_grad_(True) for t in w]
    rctx, rh, rc = vo.lstm_sequence(xr, lengths.to(DEV), *wr)
    xm = x.clone().view(B * T, 768).requires_grad_(True)
    ctx, h, c = _LSTMSeqFn.apply(xm, lengths.to(DEV), B, T, *w)
    tol = 2e-2
    assert float((ctx - rctx).abs().max()) < tol and float((h - rh).abs().max()) < tol and float((c - rc).abs().max()) < 2 * tol
    live = (torch.arange(T)[None, :] < lengths[:, None]).to(DEV)
    assert float(ctx[~live].abs().max() if (~live).any() else 0.0) == 0.0, "ctx must be exactly 0 behind a sample's end"
    g = torch.Generator().manual_seed(1)
    pc, ph, pcc = (torch.randn(t.shape, generator=g).to(DEV) for t in (rctx, rh, rc))
    ((rctx * pc).sum() + (rh * ph).sum() + (rc * pcc).sum()).backward()
    ((ctx * pc).sum() + (h * ph).sum() + (c * pcc).sum()).backward()
    grad_close("dx", xm.grad.view(B, T, 768), xr.grad, floor=1e-4)
    for name, a, r in zip(("w_ih", "w_hh", "b_ih", "b_hh"), w, wr):
        grad_close(name, a.grad, r.grad, floor=1e-4)


@pytest.mark.parametrize("space", ["vs", "tb"])
def test_agents_golden_fixture(space):
    fix = load_fixture("agents_small")
    oc = vo.OracleConfig(**fix["cfg"])
    seed, T, B, steps = fix["seed"], fix["T"], fix["B"], fix["steps"]
    P = vo.init_params(oc, seed, pretrain_heads=False)
    A, Dv, Dt = vo.init_agent_params(seed + 1, scale=fix["scale"])
    D = Dv if space == "vs" else Dt
    enc, dec = _build(oc, P, A, D, space)
    batch = _dev(synthetic.agent_batch(B, seed + 2, text_len=T, vocab=oc.vocab_size, ragged=True))
    batch["lengths"] = batch["lengths"].cpu()
    ref = fix[space]
    if space == "vs":
        trace = _dev(synthetic.viewpoint_trace(steps, B, seed + 3, candidates=6))
        loss, logits = rollout.viewpoint_rollout(enc, dec, batch, trace, steps)
    else:
        trace = _dev(synthetic.turn_trace(steps, B, seed + 3))
        loss, logits = rollout.turn_based_rollout(enc, dec, batch, trace, steps)
    assert abs(float(loss) - ref["loss"]) <= LOSS_RTOL * max(1.0, abs(ref["loss"])), (float(loss), ref["loss"])
    for t, r in enumerate(ref["logits"]):        # the reference left the loop when every episode had ended
        mine = logits[t].cpu()
        fin = torch.isfinite(r)
        assert torch.equal(torch.isfinite(mine), fin)
        _logits_close(mine, r, t)
        top2 = r.topk(2, 1).values
        decided = (top2[:, 0] - top2[:, 1]) > 2 * _logit_tol(r)[:, 0]
        assert torch.equal(mine.argmax(1)[decided], r.argmax(1)[decided])
    loss.backward()
    got = dict(enc.named_parameters())
    for k, dig in ref["enc_grads"].items():
        if k.startswith("encoder_lstm2decoder_ct"):
            continue
        digest_close(k, got[k].grad, dig)
    got = dict(dec.named_parameters())
    for k, dig in ref["dec_grads"].items():
        digest_close(k, got[k].grad, dig)
    got = dict(enc.bert.named_parameters())
    for k, dig in ref["bert_grads"].items():
        digest_close(k, got[k].grad, dig)


@pytest.mark.parametrize("space,B,feedback", [("vs", 8, "teacher"), ("tb", 4, "argmax"), ("vs", 32, "argmax")])
def test_rollout_vs_oracle_full_length(space, B, feedback):
    """T = 511 (the scripts' situation, SURVEY F6: every length 511, mask all zero), 2 encoder layers, 5 decoder steps: loss,
    per-step logits and every gradient of encoder LSTM / decoder / BERT against the fp32 oracle on the same GPU."""
    torch.backends.cuda.matmul.allow_tf32 = False
    oc = vo.OracleConfig(num_hidden_layers=2)
    steps = 5
    P = vo.init_params(oc, 41, pretrain_heads=False)
    A, Dv, Dt = vo.init_agent_params(42)
    D = Dv if space == "vs" else Dt
    enc, dec = _build(oc, P, A, D, space)
    batch = _dev(synthetic.agent_batch(B, 43))
    if space == "vs":
        trace = _dev(synthetic.viewpoint_trace(steps, B, 44))
        loss, logits = rollout.viewpoint_rollout(enc, dec, dict(batch, lengths=batch["lengths"].cpu()), trace, steps, feedback=feedback)
    else:
        trace = _dev(synthetic.turn_trace(steps, B, 44))
        loss, logits = rollout.turn_based_rollout(enc, dec, dict(batch, lengths=batch["lengths"].cpu()), trace, steps, feedback=feedback)
    loss.backward()
    Pg, Ag, Dg = (vo.clone_params(t, device=DEV) for t in (P, A, D))
    fn = vo.rollout_vs if space == "vs" else vo.rollout_tb
    rloss, rlogits = fn(Pg, oc, Ag, Dg, batch, trace, steps, feedback=feedback)
    rloss.backward()
    assert abs(float(loss) - float(rloss)) <= LOSS_RTOL * max(1.0, abs(float(rloss))), (float(loss), float(rloss))
    if feedback == "teacher":     # under argmax a near-tie may legitimately pick another action and change the later steps
        for t, r in enumerate(rlogits):
            fin = torch.isfinite(r)
            _logits_close(logits[t], r, t)
    worst = 0.0
    for k, p in list(enc.named_parameters()) + list(dec.named_parameters()):
        if k.startswith("bert."):
            ref = Pg[k].grad
        else:
            ref = (Ag if k in Ag else Dg).get(k)
            ref = None if ref is None else ref.grad
        if ref is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
            continue
        assert p.grad is not None, k
        rel, _ = grad_close(k, p.grad, ref, floor=1e-4)
        worst = max(worst, rel)
    print("worst gradient rel-L2:", worst)


def test_rollout_graph_replay_equals_eager_and_never_syncs():
    oc = vo.OracleConfig(num_hidden_layers=2)
    steps, B = 6, 4
    P = vo.init_params(oc, 51, pretrain_heads=False)
    A, Dv, _ = vo.init_agent_params(52)
    enc, dec = _build(oc, P, A, Dv, "vs")
    batch = _dev(synthetic.agent_batch(B, 53))
    batch["lengths"] = batch["lengths"]            # a DEVICE tensor of lengths: not synced on
    trace = _dev(synthetic.viewpoint_trace(steps, B, 54))
    params = list(enc.parameters()) + list(dec.parameters())

    def fwd(_model, bt):
        loss, _ = rollout.viewpoint_rollout(enc, dec, bt, trace, steps, feedback="argmax")
        return (loss,)

    loss = fwd(None, batch)[0]
    loss.backward()
    torch.cuda.synchronize()
    ref_loss = float(loss)
    ref_g = [None if p.grad is None else p.grad.clone() for p in params]
    del loss
    for p in params:
        p.grad = None
    torch.cuda.set_sync_debug_mode("error")        # the whole episode enqueues without a host <-> device sync
    try:
        l2 = fwd(None, batch)[0]
        l2.backward()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    del l2
    tensors = {k: v for k, v in batch.items() if torch.is_tensor(v)}
    step = vgraph.GraphedStep(enc, tensors, step_fn=fwd, extra_params=list(dec.parameters()))
    out = step(tensors)
    assert abs(float(out[0]) - ref_loss) < 1e-5 * max(1.0, abs(ref_loss))
    for p, r in zip(params, ref_g):
        if r is not None:
            assert float((p.grad - r).norm()) <= 1e-4 * max(float(r.norm()), 1e-9)


def test_agent_dropout_masks_match_between_forward_and_backward():
    """p = 0.5 as the scripts: the fp32 dropout regenerates its mask in backward; finite-difference-free check: the gradient of
    sum(dropout(x)) is exactly the mask times 1 / (1 - p)."""
    x = torch.randn(37, 513, device=DEV, requires_grad=True)
    y = am._DropoutFn.apply(x, 0.5, 1234, 7)
    y.sum().backward()
    keep = y != 0
    assert abs(float(keep.float().mean()) - 0.5) < 0.02
    assert torch.equal(x.grad != 0, keep) and float((x.grad[keep] - 2.0).abs().max()) < 1e-3
