"""Generates tests/golden/*.pt from the UNMODIFIED reference modules  --  TEST INFRASTRUCTURE.

Run in the build container (where /root/reference is mounted):

    python -m oracle.make_golden

The reference holds no golden vectors for this path, so these fixtures ARE
the pin: seeded parameters (``visitron_oracle.init_params``) are loaded into the reference's own
``PreTrainOscar`` / ``BertImgModelwithLocationEmbeds`` / ``SoftDotAttention`` / ``AttnDecoderLSTM``
classes, run forward+backward on seeded synthetic inputs, and the outputs are stored.  Parameters
are NOT stored (they are re-drawn from the seed); gradients are stored as (norm, sum, strided
sample) per tensor to keep every fixture small.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim, visitron_oracle as vo  # noqa: E402
from visitron_b200 import synthetic  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SAMPLE = 257  # stride used for the per-tensor gradient sample


def grad_digest(g: torch.Tensor) -> dict:
    f = g.detach().float().reshape(-1)
    return dict(norm=float(f.norm()), sum=float(f.double().sum()), sample=f[::SAMPLE][:512].clone())


def small_cfg(**kw) -> vo.OracleConfig:
    base = dict(vocab_size=2048, num_hidden_layers=2, max_position_embeddings=64, detector_classes=96,
                action_space=36)
    base.update(kw)
    return vo.OracleConfig(**base)


def gen_pretrain(ns, name, cfg, seed, batch_kw):
    P = vo.init_params(cfg, seed)
    m = ref_shim.build_pretrain(ns, cfg, P)
    m.train()  # dropout p == 0 in cfg; train() as the loop does (vs/pretrain.py:153)
    batch = synthetic.pretrain_batch(**batch_kw)
    captured = {}
    h1 = m.bert.register_forward_hook(lambda mod, i, o: captured.__setitem__("bert", o))
    h2 = m.mlmhead.register_forward_hook(lambda mod, i, o: captured.__setitem__("mlm", o))
    h3 = m.token_head.register_forward_hook(lambda mod, i, o: captured.__setitem__("tok", o))
    h4 = m.next_action.register_forward_hook(lambda mod, i, o: captured.__setitem__("act", o))
    out = m(**batch)
    out[0].backward()
    for h in (h1, h2, h3, h4):
        h.remove()
    seq, pooled = captured["bert"][0], captured["bert"][1]
    lab, tl = batch["labels"], batch["token_labels"]
    grads = {}
    for k, p in m.named_parameters():
        if k == "mlmhead.predictions.decoder.weight":
            continue  # tied: same tensor as the word embeddings
        grads[k] = grad_digest(p.grad)
    fix = dict(
        kind="pretrain", cfg=cfg.__dict__, seed=seed, batch_kw=batch_kw,
        outputs=tuple(float(o) for o in out),
        sequence_output=seq.detach().clone(), pooled_output=pooled.detach().clone(),
        mlm_logits_labelled=captured["mlm"].detach()[lab != -1][:, ::16].clone(),
        predicted_words_labelled=captured["mlm"].detach().argmax(2)[lab != -1].clone(),
        predicted_tokens_labelled=captured["tok"].detach().argmax(2)[tl != -1].clone(),
        action_scores=captured["act"].detach().clone(),
        grads=grads,
    )
    torch.save(fix, os.path.join(GOLDEN, name + ".pt"))
    print(name, fix["outputs"])


def gen_finetune(ns, name, cfg, seed, batch_kw):
    """Encoder as OscarEncoder calls it (vs/agent_models.py:267-275): text only, mask == ~uint8."""
    P = vo.init_params(cfg, seed, pretrain_heads=False)
    m = ns.BertImgModelwithLocationEmbeds(ref_shim.make_config(ns, cfg))
    m.load_state_dict({k[len("bert."):]: v for k, v in P.items()}, strict=True)
    m.train()
    batch = synthetic.finetune_encoder_batch(**batch_kw)
    seq, pooled = m(batch["input_ids"], token_type_ids=batch["token_type_ids"],
                    attention_mask=batch["attention_mask"], position_ids=None)[:2]
    g = torch.Generator().manual_seed(seed + 1)
    proj = torch.randn(seq.shape, generator=g) / seq.numel() ** 0.5
    (seq * proj).sum().backward()
    grads = {"bert." + k: grad_digest(p.grad) for k, p in m.named_parameters() if p.grad is not None}
    fix = dict(kind="finetune", cfg=cfg.__dict__, seed=seed, batch_kw=batch_kw,
               sequence_output=seq.detach().clone(), pooled_output=pooled.detach().clone(), grads=grads)
    torch.save(fix, os.path.join(GOLDEN, name + ".pt"))
    print(name, float(seq.abs().mean()))


def gen_heads(ns, name, seed):
    g = torch.Generator().manual_seed(seed)
    Hd, Fd = 512, 2052
    W = vo.init_head_params(seed, Hd, Fd)
    fix = dict(kind="heads", seed=seed)
    # --- A14 candidate scorer: vs AttnDecoderLSTM.candidate_att_layer + vs/agent.py masking + CE
    vb = synthetic.viewpoint_step_batch(8, seed)
    att = ns.vs_models.SoftDotAttention(Hd, Fd)
    att.linear_in.weight.data.copy_(W["candidate_att_layer.linear_in.weight"])
    h = vb["h_tilde"].clone().requires_grad_(True)
    cf = vb["cand_feat"].clone().requires_grad_(True)
    _, logit = att(h, cf, output_prob=False)
    C = cf.size(1)
    cand_mask = torch.arange(C)[None, :] > (vb["cand_len"] - 1)[:, None]  # utils.length2mask
    logit.masked_fill_(cand_mask, -float("inf"))
    loss = torch.nn.CrossEntropyLoss(ignore_index=-100)(logit, vb["target"])
    loss.backward()
    fix["candidate"] = dict(logit=logit.detach().clone(), loss=float(loss),
                            argmax=logit.argmax(1), dh=h.grad.clone(), dcand=grad_digest(cf.grad),
                            dw_in=grad_digest(att.linear_in.weight.grad))
    # --- A15 turn-based head: tb SoftDotAttention + decoder2action + masking + CE(ignore 7)
    tb = synthetic.turn_step_batch(6, seed, ctx_len=48)
    att2 = ns.tb_models.SoftDotAttention(Hd)
    att2.linear_in.weight.data.copy_(W["attention_layer.linear_in.weight"])
    att2.linear_out.weight.data.copy_(W["attention_layer.linear_out.weight"])
    d2a = torch.nn.Linear(Hd, 6)
    d2a.weight.data.copy_(W["decoder2action.weight"])
    d2a.bias.data.copy_(W["decoder2action.bias"])
    h2 = tb["h"].clone().requires_grad_(True)
    ctx = tb["ctx"].clone().requires_grad_(True)
    h_tilde, alpha = att2(h2, ctx, tb["ctx_mask"])
    logit2 = d2a(h_tilde)
    for i in range(logit2.size(0)):
        if bool(tb["no_forward"][i]):
            logit2[i, 4] = -float("inf")   # self.model_actions.index("forward"), tb/agent.py:71-80,316-318
    loss2 = torch.nn.CrossEntropyLoss(ignore_index=7)(logit2, tb["target"])
    loss2.backward()
    fix["turn"] = dict(h_tilde=h_tilde.detach().clone(),
                       logit=logit2.detach().clone(), loss=float(loss2), argmax=logit2.argmax(1),
                       dh=h2.grad.clone(), dctx=grad_digest(ctx.grad),
                       dw_in=grad_digest(att2.linear_in.weight.grad),
                       dw_out=grad_digest(att2.linear_out.weight.grad),
                       dw_a=d2a.weight.grad.clone(), db_a=d2a.bias.grad.clone())
    # --- A16 QA head: question_linear + BCEWithLogits(pos_weight=5, reduction none) masked mean
    dec = ns.vs_models.AttnDecoderLSTMwithClassifier(4, 64, Hd, 0.0)
    ql = dec.question_linear
    for p, k in zip((ql[0].weight, ql[0].bias, ql[2].weight, ql[2].bias),
                    ("0.weight", "0.bias", "2.weight", "2.bias")):
        p.data.copy_(W["question_linear." + k])
    hq = torch.randn(8, Hd, generator=g).requires_grad_(True)
    qt = (torch.rand(8, generator=g) < 0.4).float()
    ignore = torch.rand(8, generator=g) < 0.25
    crit = torch.nn.BCEWithLogitsLoss(reduction="none", pos_weight=torch.Tensor([5.0]))
    ql_logit = ql(hq)
    cur = crit(ql_logit.squeeze(1), qt) * ~ignore
    loss3 = cur.sum() / (len(cur) - ignore.sum())
    loss3.backward()
    fix["qa"] = dict(h=hq.detach().clone(), target=qt, ignore=ignore,
                     logit=ql_logit.detach().clone(), loss=float(loss3), dh=hq.grad.clone(),
                     dw1=grad_digest(ql[0].weight.grad), db1=ql[0].bias.grad.clone(),
                     dw2=ql[2].weight.grad.clone(), db2=ql[2].bias.grad.clone())
    torch.save(fix, os.path.join(GOLDEN, name + ".pt"))
    print(name, fix["candidate"]["loss"], fix["turn"]["loss"], fix["qa"]["loss"])


def gen_adamw(ns, name, seed):
    """The reference's AdamW + WarmupLinearSchedule (pytorch_transformers/optimization.py) for 4 steps on seeded tensors in the two
    parameter groups of vs/pretrain.py:109-127, and torch's clip_grad_norm_ as vs/agent.py:514-518 calls it."""
    g = torch.Generator().manual_seed(seed)
    shapes = [(37, 19), (5,), (1023,), (64, 33), (1,), (8200,)]
    p0 = [0.1 * torch.randn(*s, generator=g) for s in shapes]
    grads = [[torch.randn(*s, generator=g) * (0.5 + i) for s in shapes] for i in range(4)]
    params = [torch.nn.Parameter(t.clone()) for t in p0]
    groups = [dict(params=params[0::2], weight_decay=0.05), dict(params=params[1::2], weight_decay=0.0)]
    opt = ns.optimization.AdamW(groups, lr=1e-3, eps=1e-6)
    sched = ns.optimization.WarmupLinearSchedule(opt, warmup_steps=2, t_total=6)
    lrs = []
    for i in range(4):
        for p, gr in zip(params, grads[i]):
            p.grad = gr.clone()
        lrs.append([grp["lr"] for grp in opt.param_groups])
        opt.step()
        sched.step()
    clip_in = [t.clone() for t in grads[1]]
    holders = [torch.nn.Parameter(torch.zeros_like(t)) for t in clip_in]
    for h, t in zip(holders, clip_in):
        h.grad = t.clone()
    norm = torch.nn.utils.clip_grad_norm_(holders, 40.0)
    fix = dict(seed=seed, shapes=shapes, p0=p0, grads=grads, lrs=lrs, p=[p.detach().clone() for p in params],
               m=[opt.state[p]["exp_avg"].clone() for p in params], v=[opt.state[p]["exp_avg_sq"].clone() for p in params],
               clip_norm=float(norm), clipped=[h.grad.clone() for h in holders])
    torch.save(fix, os.path.join(GOLDEN, name + ".pt"))
    print(name, lrs, float(norm))


def gen_mlm_mask(name, seed):
    """The reference's own _mask_tokens on seeded token rows, with its random draws recorded (torch.bernoulli / torch.randint are
    wrapped while it runs), with and without masked_token_prediction."""
    import types as _t
    fn = ref_shim.load_mask_tokens()
    V, MASK, PAD, CLS, SEP, UNK = 2048, 103, 0, 101, 102, 100
    tok = _t.SimpleNamespace(all_special_ids=[UNK, SEP, PAD, CLS, MASK], pad_token_id=PAD, mask_token="[MASK]",
                             convert_tokens_to_ids=lambda t: MASK)
    tok_len = V

    class _Tok:
        all_special_ids = tok.all_special_ids
        pad_token_id = PAD
        mask_token = "[MASK]"

        def convert_tokens_to_ids(self, t):
            return MASK

        def __len__(self):
            return tok_len

    cases = []
    g = torch.Generator().manual_seed(seed)
    for mtp in (False, True):
        for L in (64, 511):
            ids = torch.randint(200, V, (L,), generator=g)
            ids[0] = CLS
            n_valid = int(L * 0.8)
            ids[n_valid - 1] = SEP
            ids[n_valid:] = PAD
            tc = torch.full((L,), -1, dtype=torch.long)
            pos = torch.randperm(n_valid - 2, generator=g)[: max(2, L // 16)] + 1
            tc[pos] = torch.randint(0, 96, (pos.numel(),), generator=g)
            rec = []
            ob, orr = torch.bernoulli, torch.randint

            def bern(p, *a, **k):
                out = ob(p, *a, **k)
                rec.append(out.clone().bool())
                return out

            def rint(*a, **k):
                out = orr(*a, **k)
                rec.append(out.clone())
                return out

            self_ = _t.SimpleNamespace(args=_t.SimpleNamespace(mlm_probability=0.15, masked_token_prediction=mtp), tokenizer=_Tok())
            torch.manual_seed(seed + L + int(mtp))
            torch.bernoulli, torch.randint = bern, rint
            try:
                out_ids, labels, att = fn(self_, ids.clone(), tc.clone())
            finally:
                torch.bernoulli, torch.randint = ob, orr
            assert len(rec) == 4
            cases.append(dict(mtp=mtp, ids=ids, token_classes=tc, chosen=rec[0], as_mask=rec[1], as_random=rec[2], random_words=rec[3],
                              out_ids=out_ids.clone(), labels=labels.clone(), attention=att.clone()))
    torch.save(dict(seed=seed, vocab=V, mask_id=MASK, pad_id=PAD, special=[UNK, SEP, PAD, CLS, MASK], cases=cases), os.path.join(GOLDEN, name + ".pt"))
    print(name, [(c["mtp"], int((c["labels"] != -1).sum())) for c in cases])


def gen_preprocess_item(name, seed):
    """The reference's own _preprocess_item (with its real _mask_tokens, draws recorded; the LMDB feature read replaced by seeded
    ragged arrays) for samples with fewer, exactly and more regions than max_img_seq_length."""
    import types as _t
    import numpy as np
    mask_fn, item_fn = ref_shim.load_mask_tokens(), ref_shim.load_preprocess_item()
    V, MASK, PAD, CLS, SEP, UNK, T, R, D, DL = 2048, 103, 0, 101, 102, 100, 48, 12, 70, 12

    class _Tok:
        all_special_ids = [UNK, SEP, PAD, CLS, MASK]
        pad_token_id = PAD
        mask_token = "[MASK]"

        def convert_tokens_to_ids(self, t):
            return MASK

        def __len__(self):
            return V

    g = torch.Generator().manual_seed(seed)
    cases = []
    for k, (n, mtp, nag) in enumerate([(5, True, False), (12, True, False), (17, True, True), (0, False, False), (30, False, False)]):
        ids = torch.randint(200, V, (T,), generator=g)
        ids[0], ids[T - 9], ids[T - 8:] = CLS, SEP, PAD
        tc = torch.full((T,), -1, dtype=torch.long)
        tc[torch.tensor([5, 9, 20])] = torch.randint(0, 96, (3,), generator=g)
        feats = torch.randn(n, D, generator=g).numpy().astype(np.float32)
        loc = torch.randn(n, DL, generator=g).numpy().astype(np.float32)
        rec = []
        ob, orr = torch.bernoulli, torch.randint

        def bern(p, *a, **kw):
            out = ob(p, *a, **kw)
            rec.append(out.clone().bool())
            return out

        def rint(*a, **kw):
            out = orr(*a, **kw)
            rec.append(out.clone())
            return out

        self_ = _t.SimpleNamespace(args=_t.SimpleNamespace(mlm_probability=0.15, masked_token_prediction=mtp, max_img_seq_length=R,
                                                          no_action_grounding=nag, debug=False), tokenizer=_Tok())
        self_._mask_tokens = lambda inputs, token_classes, _s=self_: mask_fn(_s, inputs, token_classes)
        self_._extract_img_features = lambda scan, vp, vi, _f=feats, _l=loc: (_f.copy(), _l.copy())
        item = dict(target_dialog_tokens_id=ids.clone(), token_classes=tc.clone(), scan="s", viewpoint="v", current_view_index=3,
                    target_rel_view_index=7 + k)
        torch.manual_seed(seed + k)
        torch.bernoulli, torch.randint = bern, rint
        try:
            out = item_fn(self_, item)
        finally:
            torch.bernoulli, torch.randint = ob, orr
        out = {kk: (v.clone() if torch.is_tensor(v) else v) for kk, v in out.items()}
        cases.append(dict(n=n, mtp=mtp, no_action_grounding=nag, ids=ids, token_classes=tc, feats=torch.from_numpy(feats), loc=torch.from_numpy(loc),
                          next_action=7 + k, chosen=rec[0], as_mask=rec[1], as_random=rec[2], random_words=rec[3], out=out))
    torch.save(dict(seed=seed, vocab=V, mask_id=MASK, pad_id=PAD, special=[UNK, SEP, PAD, CLS, MASK], R=R, cases=cases),
               os.path.join(GOLDEN, name + ".pt"))
    print(name, [(c["n"], tuple(c["out"]["img_feats"].shape), sum(c["out"]["attention_mask"])) for c in cases])



def _load(module, params):
    sd = module.state_dict()
    for k, v in params.items():
        assert k in sd and tuple(sd[k].shape) == tuple(v.shape), (k, tuple(v.shape))
        sd[k] = v.clone()
    module.load_state_dict(sd, strict=True)


def gen_agents(ns, name, seed):
    """the UNMODIFIED OscarEncoder + AttnDecoderLSTM of both action spaces (vs/agent_models.py, tb/agent_models.py)
    driven by the rollout loops of vs/agent.py:334-474 and tb/agent.py:259-392 restated over a recorded trace (agent.py itself
    needs MatterSim), teacher forcing, dropout 0: context, initial state, per-step logits, loss, gradient digests."""
    import types
    cfg = small_cfg()
    P = vo.init_params(cfg, seed, pretrain_heads=False)
    scale = 0.02      # ~ nn.Linear's default init range for a 2052-wide input: logits of O(1), as in a trained agent
    A, Dv, Dt = vo.init_agent_params(seed + 1, scale=scale)
    T, B, steps = 48, 3, 4
    args = types.SimpleNamespace(device="cpu")
    fix = dict(kind="agents", cfg=cfg.__dict__, seed=seed, T=T, B=B, steps=steps, scale=scale)
    for space in ("vs", "tb"):
        mods = ns.vs_models if space == "vs" else ns.tb_models
        bert = ns.BertImgModelwithLocationEmbeds(ref_shim.make_config(ns, cfg))
        bert.load_state_dict({k[len("bert."):]: v.clone() for k, v in P.items()}, strict=True)
        enc = mods.OscarEncoder(args, bert, 512, 512, 0.0)
        _load(enc, {k: v for k, v in A.items()})
        if space == "vs":
            dec = mods.AttnDecoderLSTM(4, 64, 512, 0.0, 2052)
            _load(dec, Dv)
        else:
            dec = mods.AttnDecoderLSTM(8, 6, 32, 512, 0.0, 2048)
            _load(dec, Dt)
        enc.train()
        dec.train()
        batch = synthetic.agent_batch(B, seed + 2, text_len=T, vocab=cfg.vocab_size, ragged=True)
        lengths, mask = batch["lengths"], batch["mask"]
        ctx, h_t, c_t = enc(batch["input_ids"], lengths, mask, token_type_ids=batch["token_type_ids"])
        out = dict(ctx=ctx.detach().clone(), decoder_init=h_t.detach().clone(), c_t=c_t.detach().clone())
        loss, logits = 0.0, []
        ended = torch.zeros(B, dtype=torch.bool)
        if space == "vs":
            tr = synthetic.viewpoint_trace(steps, B, seed + 3, candidates=6)
            crit = torch.nn.CrossEntropyLoss(ignore_index=-100)
            h1 = h_t
            for t in range(steps):
                h_t, c_t, logit, h1 = dec(tr["input_a_t"][t], tr["f_t"][t], tr["cand_feat"][t], h_t, h1, c_t, ctx, mask)
                cand_mask = torch.arange(logit.size(1))[None, :] > (tr["cand_len"][t] - 1)[:, None]     # utils.length2mask
                logit.masked_fill_(cand_mask, -float("inf"))
                target = torch.where(ended, torch.full_like(tr["target"][t], -100), tr["target"][t])
                loss = loss + crit(logit, target)
                ended = ended | (target == tr["cand_len"][t] - 1) | (target == -100)
                logits.append(logit.detach().clone())
                if bool(ended.all()):
                    break
        else:
            tr = synthetic.turn_trace(steps, B, seed + 3)
            crit = torch.nn.CrossEntropyLoss(ignore_index=7)
            a_t = torch.full((B,), 6, dtype=torch.long)
            for t in range(steps):
                h_t, c_t, alpha, logit = dec(a_t.view(-1, 1), tr["f_t"][t], h_t, c_t, ctx, mask)
                for i in range(B):
                    if bool(tr["no_forward"][t, i]):
                        logit[i, 4] = -float("inf")          # model_actions.index("forward")
                target = torch.where(ended, torch.full_like(tr["target"][t], 7), tr["target"][t])
                loss = loss + crit(logit, target)
                a_t = target
                ended = ended | (a_t == 5)
                logits.append(logit.detach().clone())
                if bool(ended.all()):
                    break
        loss = loss / steps
        loss.backward()
        out.update(loss=float(loss), logits=logits,
                   enc_grads={k: grad_digest(p.grad) for k, p in enc.named_parameters() if p.grad is not None and not k.startswith("bert.")},
                   dec_grads={k: grad_digest(p.grad) for k, p in dec.named_parameters() if p.grad is not None},
                   bert_grads={k: grad_digest(p.grad) for k, p in enc.bert.named_parameters()
                               if p.grad is not None and ("layer.0.attention.self.query" in k or "word_embeddings" in k or "layer.1.output.dense.weight" in k)})
        fix[space] = out
        print(name, space, float(loss))
    torch.save(fix, os.path.join(GOLDEN, name + ".pt"))


def main():
    if not ref_shim.available():
        raise SystemExit("reference tree not mounted; fixtures can only be generated in the build container")
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ns = ref_shim.load()
    bk = dict(batch=2, seed=88, text_len=40, regions=24, img_dim=2054, vocab=2048, detector_classes=96)
    gen_pretrain(ns, "pretrain_small", small_cfg(), 7, bk)
    gen_pretrain(ns, "pretrain_small_imgln", small_cfg(use_img_layernorm=1), 8, dict(bk, seed=89))
    gen_finetune(ns, "finetune_small", small_cfg(), 9, dict(batch=2, seed=90, text_len=48, vocab=2048))
    gen_heads(ns, "heads_small", 11)
    gen_agents(ns, "agents_small", 23)
    gen_adamw(ns, "adamw_small", 13)
    gen_mlm_mask("mlm_mask_small", 17)
    gen_preprocess_item("preprocess_item_small", 19)


if __name__ == "__main__":
    main()
