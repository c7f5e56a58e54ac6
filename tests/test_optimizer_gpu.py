"""The step after backward on the B200 kernels, through the C ABI: multi-tensor AdamW, the warm-up
schedule and gradient-norm clipping against (a) the golden fixture the UNMODIFIED reference optimizer produced
(tests/golden/adamw_small.pt) and (b) the oracle restatement at the full size of the pre-training model.
fp32 end to end; tolerance: a few ulps (2e-6 relative, 2e-7 absolute), the clipping norm 1e-6 relative."""
import pytest
import torch

from oracle import visitron_oracle as vo
from parity_common import load_fixture
from visitron_b200 import lib, optimization as vopt

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, rtol=2e-6, atol=2e-7):
    torch.testing.assert_close(a.detach().cpu(), b.detach().cpu(), rtol=rtol, atol=atol)


def test_adamw_schedule_golden():
    fix = load_fixture("adamw_small")
    params = [torch.nn.Parameter(t.clone().to(DEV)) for t in fix["p0"]]
    groups = [dict(params=params[0::2], weight_decay=0.05), dict(params=params[1::2], weight_decay=0.0)]   # vs/pretrain.py:109-127
    opt = vopt.AdamW(groups, lr=1e-3, eps=1e-6)
    sched = vopt.WarmupLinearSchedule(opt, warmup_steps=2, t_total=6)
    for i in range(4):
        for p, g in zip(params, fix["grads"][i]):
            p.grad = g.clone().to(DEV)
        assert [grp["lr"] for grp in opt.param_groups] == pytest.approx(fix["lrs"][i], abs=1e-12)
        n0 = lib.launch_count()
        v0 = params[0]._version
        opt.step()
        assert lib.launch_count() - n0 == 1, "every parameter of every group in ONE launch"
        assert params[0]._version > v0, "autograd must see the in-place update"
        sched.step()
    for i, p in enumerate(params):
        _close(p, fix["p"][i])
        _close(opt.state[p]["exp_avg"], fix["m"][i])
        _close(opt.state[p]["exp_avg_sq"], fix["v"][i])
        assert opt.state[p]["step"] == 4


def test_clip_grad_norm_golden():
    fix = load_fixture("adamw_small")
    holders = [torch.nn.Parameter(torch.zeros_like(t).to(DEV)) for t in fix["grads"][1]]
    for h, t in zip(holders, fix["grads"][1]):
        h.grad = t.clone().to(DEV)
    norm = vopt.clip_grad_norm_(holders, 40.0)
    assert abs(float(norm) - fix["clip_norm"]) < 1e-6 * fix["clip_norm"]
    for h, t in zip(holders, fix["clipped"]):
        _close(h.grad, t, rtol=1e-5)
    # below the bound nothing changes, and the call is deterministic
    before = [h.grad.clone() for h in holders]
    n1 = vopt.clip_grad_norm_(holders, 1e6)
    n2 = vopt.clip_grad_norm_(holders, 1e6)
    assert float(n1) == float(n2)
    for h, t in zip(holders, before):
        assert torch.equal(h.grad, t)


def test_clip_folded_into_step_equals_clip_then_step():
    g = torch.Generator().manual_seed(3)
    p0 = [torch.randn(1000, generator=g), torch.randn(33, 7, generator=g)]
    gr = [10 * torch.randn(1000, generator=g), 10 * torch.randn(33, 7, generator=g)]
    res = []
    for folded in (False, True):
        ps = [torch.nn.Parameter(t.clone().to(DEV)) for t in p0]
        for p, t in zip(ps, gr):
            p.grad = t.clone().to(DEV)
        opt = vopt.AdamW(ps, lr=1e-2, weight_decay=0.01)
        vopt.clip_grad_norm_(ps, 5.0, optimizer=opt if folded else None)
        if folded:
            assert torch.equal(ps[0].grad.cpu(), gr[0]), "folded clipping leaves the gradients alone"
        opt.step()
        res.append([p.detach().clone() for p in ps])
    for a, b in zip(*res):
        _close(a, b, rtol=1e-6, atol=1e-7)


def test_adamw_full_model_against_oracle_and_refreshes_operands():
    """bert-base pre-training model (113 M parameters in the flat buffer): one fused step == the oracle's per-tensor update,
    and the next forward runs on the NEW weights (the bf16 operand cache notices the version bump)."""
    from visitron_b200 import modeling, synthetic
    from visitron_b200.pretrained import BertConfig
    cfg = BertConfig(30522, num_hidden_layers=2, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    cfg.img_feature_dim, cfg.detector_classes, cfg.action_space = 2054, 1601, 36
    torch.manual_seed(0)
    model = modeling.PreTrainOscar(cfg).to(DEV).train()
    batch = {k: v.to(DEV) for k, v in synthetic.pretrain_batch(2, 5, text_len=64, regions=16).items()}
    no_decay = ["bias", "LayerNorm.weight"]
    named = list(model.named_parameters())
    groups = [dict(params=[p for n, p in named if not any(nd in n for nd in no_decay)], weight_decay=0.01),
              dict(params=[p for n, p in named if any(nd in n for nd in no_decay)], weight_decay=0.0)]
    opt = vopt.AdamW(groups, lr=5e-5, eps=1e-8)
    out0 = model(**batch)
    out0[0].backward()
    before = {n: (p.detach().clone(), p.grad.detach().clone()) for n, p in named if p.grad is not None}
    n0 = lib.launch_count()
    opt.step()
    assert lib.launch_count() - n0 == 1
    for n in ("bert.encoder.layer.1.output.dense.weight", "bert.embeddings.word_embeddings.weight", "bert.encoder.layer.0.attention.self.key.bias",
              "bert.embeddings.LayerNorm.weight", "token_head.0.weight"):
        p, g = before[n]
        wd = 0.0 if any(nd in n for nd in no_decay) else 0.01
        want, _, _ = vo.adamw_update(p, g, torch.zeros_like(p), torch.zeros_like(p), 1, 5e-5, eps=1e-8, weight_decay=wd)
        _close(dict(named)[n], want)
    with torch.no_grad():
        out1 = model(**batch)
    assert float(out1[0]) != float(out0[0]), "the forward after step() must use the updated weights"
