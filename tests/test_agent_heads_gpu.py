"""Fine-tuning heads on the B200 kernels, through the C ABI, against (a) the golden fixtures the
UNMODIFIED reference modules produced (tests/golden/heads_small.pt) and (b) the oracle at the full shapes of the
fine-tuning configs.  fp32 end to end: tolerance 1e-4 relative / 1e-5 absolute on the fixtures, 5e-4 / 5e-5 on the
full-size random cases (their softmax logits have a standard deviation of ~25, so fp32 summation order shows), argmax exact."""
import pytest
import torch

from oracle import visitron_oracle as vo
from parity_common import load_fixture
from test_oracle import _check_digest
from visitron_b200 import agent_heads as ah, synthetic

pytestmark = pytest.mark.gpu
DEV = "cuda"
RT, AT = 1e-4, 1e-5


def _dev(d):
    return {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in d.items()}


def _leaf(t):
    return t.detach().clone().to(DEV).requires_grad_(True)


def _close(a, b, rtol=RT, atol=AT, msg=None):
    torch.testing.assert_close(a.detach().cpu(), b.detach().cpu(), rtol=rtol, atol=atol, msg=msg)


def test_candidate_scorer_golden():
    fix = load_fixture("heads_small")
    W = vo.init_head_params(fix["seed"])
    vb = _dev(synthetic.viewpoint_step_batch(8, fix["seed"]))
    w_in, h, cf = _leaf(W["candidate_att_layer.linear_in.weight"]), _leaf(vb["h_tilde"]), _leaf(vb["cand_feat"])
    logit, loss, am = ah.candidate_scores_ce(w_in, h, cf, vb["cand_len"], vb["target"])
    loss.backward()
    c = fix["candidate"]
    _close(logit, c["logit"])
    assert abs(float(loss) - c["loss"]) < 1e-5 * max(1, c["loss"])
    assert torch.equal(am.cpu(), c["argmax"])
    _close(h.grad, c["dh"], atol=1e-6)
    _check_digest("dcand", cf.grad.cpu(), c["dcand"])
    _check_digest("dw_in", w_in.grad.cpu(), c["dw_in"])


def test_turn_based_head_golden():
    fix = load_fixture("heads_small")
    W = vo.init_head_params(fix["seed"])
    tb = _dev(synthetic.turn_step_batch(6, fix["seed"], ctx_len=48))
    ws = [_leaf(W[k]) for k in ("attention_layer.linear_in.weight", "attention_layer.linear_out.weight", "decoder2action.weight",
                                "decoder2action.bias")]
    h, ctx = _leaf(tb["h"]), _leaf(tb["ctx"])
    ht, logit, loss, am = ah.turn_based_head(*ws, h, ctx, tb["ctx_mask"], tb["no_forward"], tb["target"])
    loss.backward()
    t = fix["turn"]
    _close(ht, t["h_tilde"])
    _close(logit, t["logit"])
    assert abs(float(loss) - t["loss"]) < 1e-5 and torch.equal(am.cpu(), t["argmax"])
    _close(h.grad, t["dh"], atol=1e-6)
    _check_digest("dctx", ctx.grad.cpu(), t["dctx"])
    _check_digest("dw_in", ws[0].grad.cpu(), t["dw_in"])
    _check_digest("dw_out", ws[1].grad.cpu(), t["dw_out"])
    _close(ws[2].grad, t["dw_a"], atol=1e-6)
    _close(ws[3].grad, t["db_a"], atol=1e-6)


def test_qa_head_golden():
    fix = load_fixture("heads_small")
    W = vo.init_head_params(fix["seed"])
    q = fix["qa"]
    qs = [_leaf(W["question_linear." + k]) for k in ("0.weight", "0.bias", "2.weight", "2.bias")]
    hq = _leaf(q["h"])
    logit, loss = ah.qa_head(*qs, hq, q["target"].to(DEV), (~q["ignore"]).to(DEV))
    loss.backward()
    _close(logit, q["logit"])
    assert abs(float(loss) - q["loss"]) < 1e-5
    _close(hq.grad, q["dh"], atol=1e-6)
    _check_digest("dw1", qs[0].grad.cpu(), q["dw1"])
    _close(qs[3].grad, q["db2"], atol=1e-6)


@pytest.mark.parametrize("B,C", [(8, 12), (64, 15), (1, 2)])
def test_candidate_scorer_vs_oracle_full_width(B, C):
    """vs fine-tuning step: h~ [B,512], candidates [B,C,2052] (2048 + 4 angle features, vs/agent.py:124)."""
    W = vo.init_head_params(3)
    vb = synthetic.viewpoint_step_batch(B, 5, candidates=C)
    ref_in = [W["candidate_att_layer.linear_in.weight"].clone().requires_grad_(True), vb["h_tilde"].clone().requires_grad_(True),
              vb["cand_feat"].clone().requires_grad_(True)]
    rl, rloss, ram = vo.candidate_scores_ce(*ref_in, vb["cand_len"], vb["target"])
    mine_in = [_leaf(t) for t in ref_in]
    ml, mloss, mam = ah.candidate_scores_ce(*mine_in, vb["cand_len"].to(DEV), vb["target"].to(DEV))
    _close(ml, rl)
    assert torch.equal(mam.cpu(), ram)
    if bool((vb["target"] != -100).any()):
        rloss.backward()
        mloss.backward()
        assert abs(float(mloss) - float(rloss)) < 1e-5 * max(1.0, abs(float(rloss)))
        for a, b, n in zip(mine_in, ref_in, ("dw_in", "dh", "dcand")):
            _close(a.grad, b.grad, rtol=1e-3, atol=1e-6, msg=n)
    else:
        assert torch.isnan(mloss) and torch.isnan(rloss)  # every episode ended: CrossEntropyLoss gives 0/0


@pytest.mark.parametrize("B,L", [(4, 511), (32, 511), (3, 80)])
def test_turn_based_head_vs_oracle_full_context(B, L):
    """tb fine-tuning step: ctx [B,511,512] from the encoder LSTM, 6 actions, ignore_index 7."""
    W = vo.init_head_params(4)
    tb = synthetic.turn_step_batch(B, 9, ctx_len=L)
    keys = ("attention_layer.linear_in.weight", "attention_layer.linear_out.weight", "decoder2action.weight", "decoder2action.bias")
    rw = [W[k].clone().requires_grad_(True) for k in keys]
    rh, rc = tb["h"].clone().requires_grad_(True), tb["ctx"].clone().requires_grad_(True)
    rht, rl, rloss, ram = vo.turn_based_head(*rw, rh, rc, tb["ctx_mask"], tb["no_forward"], tb["target"])
    rloss.backward()
    mw = [_leaf(t) for t in rw]
    mh, mc = _leaf(rh), _leaf(rc)
    mht, ml, mloss, mam = ah.turn_based_head(*mw, mh, mc, tb["ctx_mask"].to(DEV), tb["no_forward"].to(DEV), tb["target"].to(DEV))
    mloss.backward()
    _close(mht, rht, rtol=5e-4, atol=5e-5)
    _close(ml, rl, rtol=5e-4, atol=5e-5)
    assert torch.equal(mam.cpu(), ram) and abs(float(mloss) - float(rloss)) < 1e-4 * max(1.0, abs(float(rloss)))
    _close(mh.grad, rh.grad, rtol=1e-3, atol=1e-5)
    _close(mc.grad, rc.grad, rtol=1e-3, atol=1e-6)
    for a, b, k in zip(mw, rw, keys):
        _close(a.grad, b.grad, rtol=1e-3, atol=1e-5, msg=k)


def test_soft_dot_attention_module_is_a_drop_in():
    """Same parameters and forward() variants as the reference class (vs/agent_models.py:313-357); the view-feature
    attention of the decoder (feat_att_layer, output_tilde=False over [B,36,2052]) and the masked context attention."""
    torch.manual_seed(0)
    B = 5
    att = ah.SoftDotAttention(512, 2052).to(DEV)
    assert set(att.state_dict()) == {"linear_in.weight", "linear_out.weight"}
    assert att.linear_in.weight.shape == (2052, 512) and att.linear_out.weight.shape == (512, 2564)
    h = torch.randn(B, 512, device=DEV, requires_grad=True)
    feat = 0.3 * torch.randn(B, 36, 2052, device=DEV)
    wc, attn = att(h, feat, output_tilde=False)
    w_in, w_out = att.linear_in.weight.detach().cpu(), att.linear_out.weight.detach().cpu()
    rh = h.detach().cpu().requires_grad_(True)
    rwc, rattn = vo.soft_dot_attention(w_in, w_out, rh, feat.cpu(), output_tilde=False)
    _close(wc, rwc, rtol=5e-4, atol=5e-5)
    _close(attn, rattn, rtol=5e-4, atol=5e-5)
    (wc.square().sum() + attn[:, 0].sum()).backward()
    (rwc.square().sum() + rattn[:, 0].sum()).backward()
    _close(h.grad, rh.grad, rtol=2e-3, atol=1e-4)
    # tb flavour: SoftDotAttention(dim), masked, output_tilde + probabilities; masked positions are exactly zero / -inf
    att2 = ah.SoftDotAttention(512).to(DEV)
    ctx = torch.randn(B, 100, 512, device=DEV, requires_grad=True)
    mask = torch.arange(100, device=DEV)[None, :] >= torch.tensor([100, 60, 1, 99, 50], device=DEV)[:, None]
    ht, alpha = att2(h, ctx, mask)
    rht, ralpha = vo.soft_dot_attention(att2.linear_in.weight.detach().cpu(), att2.linear_out.weight.detach().cpu(), h.detach().cpu(),
                                        ctx.detach().cpu(), mask.cpu())
    _close(ht, rht, rtol=5e-4, atol=5e-5)
    _close(alpha, ralpha, rtol=5e-4, atol=5e-5)
    assert float(alpha[mask].abs().max()) == 0.0
    _, logit = att2(h, ctx, mask, output_prob=False)
    assert bool(torch.isinf(logit[mask]).all()) and bool(torch.isfinite(logit[~mask]).all())


def test_masked_cross_entropy_edges():
    logit = torch.randn(6, 9, device=DEV, requires_grad=True)
    target = torch.tensor([0, 3, -100, 8, 1, 2], device=DEV)
    lens = torch.tensor([9, 4, 2, 9, 2, 3], device=DEV)
    loss, masked, am = ah.masked_cross_entropy(logit, target, valid_len=lens)
    ref = logit.detach().cpu().clone().requires_grad_(True)
    rm = ref.masked_fill(torch.arange(9)[None, :] >= lens.cpu()[:, None], -float("inf"))
    rloss = torch.nn.functional.cross_entropy(rm, target.cpu(), ignore_index=-100)
    _close(loss, rloss)
    assert torch.equal(masked.cpu(), rm.detach()) and torch.equal(am.cpu(), rm.argmax(1))
    (2.0 * loss).backward()
    (2.0 * rloss).backward()
    _close(logit.grad, ref.grad, atol=1e-7)
    # ties break to the lowest index, like torch.argmax
    flat = torch.zeros(2, 40, device=DEV)
    _, _, am2 = ah.masked_cross_entropy(flat, torch.zeros(2, dtype=torch.long, device=DEV))
    assert am2.tolist() == [0, 0]
    # every row ignored -> nan (0/0), like CrossEntropyLoss
    l3, _, _ = ah.masked_cross_entropy(flat, torch.full((2,), 7, device=DEV), ignore_index=7)
    assert bool(torch.isnan(l3))


def test_bce_masked_mean_edges():
    q = torch.randn(7, 1, device=DEV, requires_grad=True)
    y = torch.tensor([1, 0, 0, 1, 0, 1, 0], device=DEV)
    valid = torch.tensor([1, 1, 0, 1, 0, 1, 1], dtype=torch.bool, device=DEV)
    loss, per = ah.bce_with_logits_masked_mean(q, y, valid, 5.0)
    rq = q.detach().cpu().clone().requires_grad_(True)
    rper = torch.nn.functional.binary_cross_entropy_with_logits(rq.squeeze(1), y.cpu().float(), reduction="none", pos_weight=torch.tensor(5.0))
    rloss = (rper * valid.cpu()).sum() / valid.sum().cpu()
    _close(loss, rloss)
    _close(per.squeeze(1), rper)
    loss.backward()
    rloss.backward()
    _close(q.grad, rq.grad, atol=1e-7)
    l0, _ = ah.bce_with_logits_masked_mean(q.detach(), y, torch.zeros(7, dtype=torch.bool, device=DEV))
    assert float(l0) == 0.0  # nothing valid: the reference multiplies by 0 (vs/classifier/agent.py:495-497)


def test_heads_refuse_cpu_tensors():
    from visitron_b200 import lib
    with pytest.raises(lib.VtError):
        ah.small_linear(torch.randn(2, 8), torch.randn(4, 8))
