"""The oracle (oracle/visitron_oracle.py) against (a) the committed golden fixtures produced by the
UNMODIFIED reference modules and (b), when /root/reference is mounted, the live reference on other
seeds / configurations.  CPU only."""
import os

import pytest
import torch

from oracle import ref_shim, visitron_oracle as vo
from oracle.make_golden import SAMPLE
from visitron_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def _check_digest(name, g, dig, rtol=2e-4):
    f = g.detach().float().reshape(-1)
    assert abs(float(f.norm()) - dig["norm"]) <= rtol * max(dig["norm"], 1e-6), name
    torch.testing.assert_close(f[::SAMPLE][:512], dig["sample"], rtol=1e-3, atol=1e-6 + 1e-4 * dig["norm"] / f.numel() ** 0.5, msg=name)


@pytest.mark.parametrize("name", ["pretrain_small", "pretrain_small_imgln"])
def test_pretrain_golden(name):
    fix = _load(name)
    cfg = vo.OracleConfig(**fix["cfg"])
    P = vo.clone_params(vo.init_params(cfg, fix["seed"]))
    batch = synthetic.pretrain_batch(**fix["batch_kw"])
    out, aux = vo.pretrain_forward(P, cfg, return_aux=True, **batch)
    out[0].backward()
    for a, b in zip(out, fix["outputs"]):
        assert abs(float(a) - b) <= 1e-5 * max(1.0, abs(b))
    torch.testing.assert_close(aux["sequence_output"], fix["sequence_output"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(aux["pooled_output"], fix["pooled_output"], rtol=1e-4, atol=1e-5)
    lab, tl = batch["labels"], batch["token_labels"]
    torch.testing.assert_close(aux["prediction_scores"][lab != -1][:, ::16], fix["mlm_logits_labelled"], rtol=1e-4, atol=1e-5)
    assert torch.equal(aux["prediction_scores"].argmax(2)[lab != -1], fix["predicted_words_labelled"])
    assert torch.equal(aux["token_prob"].argmax(2)[tl != -1], fix["predicted_tokens_labelled"])
    torch.testing.assert_close(aux["action_scores"], fix["action_scores"], rtol=1e-4, atol=1e-5)
    assert set(fix["grads"]) == set(P)
    for k, dig in fix["grads"].items():
        _check_digest(k, P[k].grad, dig)


def test_finetune_golden():
    fix = _load("finetune_small")
    cfg = vo.OracleConfig(**fix["cfg"])
    P = vo.clone_params(vo.init_params(cfg, fix["seed"], pretrain_heads=False))
    batch = synthetic.finetune_encoder_batch(**fix["batch_kw"])
    assert batch["attention_mask"].dtype == torch.uint8 and int(batch["attention_mask"].min()) == 255  # F6
    seq, pooled = vo.bert_img_forward(P, cfg, **batch)
    torch.testing.assert_close(seq, fix["sequence_output"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(pooled, fix["pooled_output"], rtol=1e-4, atol=1e-5)
    g = torch.Generator().manual_seed(fix["seed"] + 1)
    proj = torch.randn(seq.shape, generator=g) / seq.numel() ** 0.5
    (seq * proj).sum().backward()
    for k, dig in fix["grads"].items():
        _check_digest(k, P[k].grad, dig)
    # parameters the fine-tune call leaves without gradient (needs find_unused_parameters in the reference)
    for k in ("bert.img_embedding.weight", "bert.location_embeds.weight", "bert.pooler.dense.weight"):
        assert k not in fix["grads"] and P[k].grad is None


def test_heads_golden():
    fix = _load("heads_small")
    W = vo.init_head_params(fix["seed"])
    # A14
    vb = synthetic.viewpoint_step_batch(8, fix["seed"])
    w_in = W["candidate_att_layer.linear_in.weight"].clone().requires_grad_(True)
    h = vb["h_tilde"].clone().requires_grad_(True)
    cf = vb["cand_feat"].clone().requires_grad_(True)
    logit, loss, am = vo.candidate_scores_ce(w_in, h, cf, vb["cand_len"], vb["target"])
    loss.backward()
    c = fix["candidate"]
    torch.testing.assert_close(logit, c["logit"], rtol=1e-4, atol=1e-5)
    assert abs(float(loss) - c["loss"]) < 1e-5 * max(1, c["loss"]) and torch.equal(am, c["argmax"])
    torch.testing.assert_close(h.grad, c["dh"], rtol=1e-4, atol=1e-6)
    _check_digest("dcand", cf.grad, c["dcand"])
    _check_digest("dw_in", w_in.grad, c["dw_in"])
    # A15
    tb = synthetic.turn_step_batch(6, fix["seed"], ctx_len=48)
    ws = [W[k].clone().requires_grad_(True) for k in ("attention_layer.linear_in.weight", "attention_layer.linear_out.weight", "decoder2action.weight", "decoder2action.bias")]
    h2 = tb["h"].clone().requires_grad_(True)
    ctx = tb["ctx"].clone().requires_grad_(True)
    ht, logit2, loss2, am2 = vo.turn_based_head(*ws, h2, ctx, tb["ctx_mask"], tb["no_forward"], tb["target"])
    loss2.backward()
    t = fix["turn"]
    torch.testing.assert_close(ht, t["h_tilde"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(logit2, t["logit"], rtol=1e-4, atol=1e-5)
    assert abs(float(loss2) - t["loss"]) < 1e-5 and torch.equal(am2, t["argmax"])
    torch.testing.assert_close(h2.grad, t["dh"], rtol=1e-4, atol=1e-6)
    _check_digest("dctx", ctx.grad, t["dctx"])
    _check_digest("dw_in", ws[0].grad, t["dw_in"])
    _check_digest("dw_out", ws[1].grad, t["dw_out"])
    torch.testing.assert_close(ws[2].grad, t["dw_a"], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(ws[3].grad, t["db_a"], rtol=1e-4, atol=1e-6)
    # A16
    q = fix["qa"]
    qs = [W["question_linear." + k].clone().requires_grad_(True) for k in ("0.weight", "0.bias", "2.weight", "2.bias")]
    hq = q["h"].clone().requires_grad_(True)
    logit3, loss3 = vo.qa_head(*qs, hq, q["target"], ~q["ignore"])
    loss3.backward()
    torch.testing.assert_close(logit3, q["logit"], rtol=1e-4, atol=1e-5)
    assert abs(float(loss3) - q["loss"]) < 1e-5
    torch.testing.assert_close(hq.grad, q["dh"], rtol=1e-4, atol=1e-6)
    _check_digest("dw1", qs[0].grad, q["dw1"])
    torch.testing.assert_close(qs[3].grad, q["db2"], rtol=1e-4, atol=1e-6)


@pytest.mark.skipif(not ref_shim.available(), reason="reference tree not mounted (GPU box)")
@pytest.mark.parametrize("seed,imgln,layers", [(21, 0, 1), (22, 1, 3)])
def test_oracle_vs_live_reference(seed, imgln, layers):
    """Full-tensor comparison (all gradients, not digests) on configurations the fixtures do not hold."""
    ns = ref_shim.load()
    cfg = vo.OracleConfig(vocab_size=1500, num_hidden_layers=layers, max_position_embeddings=40,
                          detector_classes=50, use_img_layernorm=imgln)
    P0 = vo.init_params(cfg, seed)
    m = ref_shim.build_pretrain(ns, cfg, P0)
    m.train()
    batch = synthetic.pretrain_batch(3, seed, text_len=33, regions=16, vocab=1500, detector_classes=50)
    ref = m(**batch)
    ref[0].backward()
    P = vo.clone_params(P0)
    out = vo.pretrain_forward(P, cfg, **batch)
    out[0].backward()
    for a, b in zip(out, ref):
        assert abs(float(a) - float(b)) <= 1e-5 * max(1.0, abs(float(b)))
    for k, p in m.named_parameters():
        if k == "mlmhead.predictions.decoder.weight":
            continue
        torch.testing.assert_close(P[k].grad, p.grad, rtol=2e-3, atol=1e-6, msg=k)


def test_mask_quantisation_f6():
    """~uint8 masks give bias (1-255)*-1e4 = +2.54e6: softmax is shift invariant but fp32 scores are
    quantised to 0.25 -- the oracle must reproduce that, not an exact unmasked softmax."""
    ext = vo.extended_mask(torch.full((1, 4), 255, dtype=torch.uint8), torch.float32)
    assert float(ext[0, 0, 0, 0]) == 2540000.0
    s = torch.tensor([0.1, 0.2, 0.3, 0.45]) + ext[0, 0, 0]
    assert torch.equal(s - 2540000.0, torch.tensor([0.0, 0.25, 0.25, 0.5]))


def test_adamw_oracle_matches_reference_fixture():
    """oracle.adamw_update / warmup_linear / clip_grad_norm against the reference's AdamW + WarmupLinearSchedule run
    (tests/golden/adamw_small.pt, generated by oracle/make_golden.py from pytorch_transformers/optimization.py)."""
    fix = torch.load(os.path.join(GOLDEN, "adamw_small.pt"), weights_only=False)
    n = len(fix["shapes"])
    p = [t.clone() for t in fix["p0"]]
    m = [torch.zeros_like(t) for t in p]
    v = [torch.zeros_like(t) for t in p]
    for step in range(4):
        lr = 1e-3 * vo.warmup_linear(step, 2, 6)
        assert abs(lr - fix["lrs"][step][0]) < 1e-12
        for i in range(n):
            wd = 0.05 if i % 2 == 0 else 0.0   # groups: params[0::2] decay, params[1::2] none
            p[i], m[i], v[i] = vo.adamw_update(p[i], fix["grads"][step][i], m[i], v[i], step + 1, lr, eps=1e-6, weight_decay=wd)
    for i in range(n):
        # fp32 rounding only (the restatement is functional, the reference updates in place): a few ulps
        torch.testing.assert_close(p[i], fix["p"][i], rtol=2e-6, atol=2e-7)
        torch.testing.assert_close(m[i], fix["m"][i], rtol=2e-6, atol=2e-7)
        torch.testing.assert_close(v[i], fix["v"][i], rtol=2e-6, atol=2e-7)
    norm, clipped = vo.clip_grad_norm(fix["grads"][1], 40.0)
    assert abs(float(norm) - fix["clip_norm"]) < 1e-4 * fix["clip_norm"]
    for a, b in zip(clipped, fix["clipped"]):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-7)


def test_mlm_mask_oracle_matches_reference_fixture():
    """oracle.mlm_mask against the reference's own PretrainDataset._mask_tokens (tests/golden/mlm_mask_small.pt: its inputs, the
    bernoulli / randint draws it made, and its outputs; generated by oracle/make_golden.py), with and without
    masked_token_prediction.  Integer work: bit-exact."""
    fix = torch.load(os.path.join(GOLDEN, "mlm_mask_small.pt"), weights_only=False)
    assert len(fix["cases"]) == 4
    for c in fix["cases"]:
        special = torch.zeros(fix["vocab"], dtype=torch.bool)
        special[torch.tensor(fix["special"])] = True
        # the reference draws bernoulli(0) on special tokens, so `chosen` is already clear there; the oracle takes raw draws
        ids, labels, att = vo.mlm_mask(c["ids"], c["token_classes"] if c["mtp"] else None, special[c["ids"]], c["chosen"], c["as_mask"],
                                       c["as_random"], c["random_words"], fix["mask_id"], fix["pad_id"])
        assert torch.equal(ids, c["out_ids"]) and torch.equal(labels, c["labels"]) and torch.equal(att, c["attention"])
        if c["mtp"]:
            forced = c["token_classes"] != -1
            assert bool((ids[forced] == fix["mask_id"]).all()) and bool((labels[forced] == -1).all())


def test_preprocess_item_oracle_matches_reference_fixture():
    """oracle.mlm_mask + oracle.preprocess_item against the reference's own PretrainDataset._preprocess_item
    (tests/golden/preprocess_item_small.pt): samples with fewer / exactly / more regions than max_img_seq_length and none at all,
    with and without masked_token_prediction / no_action_grounding.  Bit-exact."""
    fix = torch.load(os.path.join(GOLDEN, "preprocess_item_small.pt"), weights_only=False)
    special = torch.zeros(fix["vocab"], dtype=torch.bool)
    special[torch.tensor(fix["special"])] = True
    for c in fix["cases"]:
        masked = vo.mlm_mask(c["ids"], c["token_classes"] if c["mtp"] else None, special[c["ids"]], c["chosen"], c["as_mask"], c["as_random"],
                             c["random_words"], fix["mask_id"], fix["pad_id"])
        got = vo.preprocess_item(masked, c["token_classes"], c["feats"], c["loc"], c["next_action"], fix["R"], c["mtp"], c["no_action_grounding"])
        want = c["out"]
        assert set(got) == set(want)
        for k, v in want.items():
            if torch.is_tensor(v):
                assert torch.equal(got[k], v), (c["n"], k)
            else:
                assert got[k] == v or (k == "attention_mask" and [int(a) for a in v] == got[k]), (c["n"], k)


@pytest.mark.parametrize("space", ["vs", "tb"])
def test_agents_golden(space):
    """the oracle's OscarEncoder LSTM, decoder steps and rollout loops against the fixture produced by the
    UNMODIFIED OscarEncoder / AttnDecoderLSTM of tasks/{viewpoint_select,turn_based}/agent_models.py (teacher forcing, ragged
    lengths through pack_padded_sequence, an episode that ends early)."""
    fix = _load("agents_small")
    cfg = vo.OracleConfig(**fix["cfg"])
    seed, T, B, steps = fix["seed"], fix["T"], fix["B"], fix["steps"]
    P = vo.clone_params(vo.init_params(cfg, seed, pretrain_heads=False))
    A, Dv, Dt = vo.init_agent_params(seed + 1, scale=fix["scale"])
    A, D = vo.clone_params(A), vo.clone_params(Dv if space == "vs" else Dt)
    batch = synthetic.agent_batch(B, seed + 2, text_len=T, vocab=cfg.vocab_size, ragged=True)
    ref = fix[space]
    ctx, h0, c0 = vo.oscar_encoder_forward(P, cfg, A, batch["input_ids"], batch["lengths"], batch["mask"], batch["token_type_ids"])
    torch.testing.assert_close(ctx, ref["ctx"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(h0, ref["decoder_init"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(c0, ref["c_t"], rtol=1e-4, atol=1e-5)
    if space == "vs":
        trace = synthetic.viewpoint_trace(steps, B, seed + 3, candidates=6)
        loss, logits = vo.rollout_vs(P, cfg, A, D, batch, trace, steps)
    else:
        trace = synthetic.turn_trace(steps, B, seed + 3)
        loss, logits = vo.rollout_tb(P, cfg, A, D, batch, trace, steps)
    assert len(logits) == len(ref["logits"])
    for a, b in zip(logits, ref["logits"]):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)
    assert abs(float(loss) - ref["loss"]) <= 1e-5 * max(1.0, abs(ref["loss"]))
    loss.backward()
    for k, dig in ref["enc_grads"].items():
        if k in ("encoder_lstm2decoder_ct.weight", "encoder_lstm2decoder_ct.bias"):
            continue
        _check_digest(k, A[k].grad, dig)
    for k, dig in ref["dec_grads"].items():
        _check_digest(k, D[k].grad, dig)
    for k, dig in ref["bert_grads"].items():
        _check_digest(k, P["bert." + k].grad, dig)
