"""Shared pieces of the model-level parity tests (CPU host-logic suite and the B200 suite)."""
import os

import torch

from oracle import visitron_oracle as vo
from oracle.make_golden import SAMPLE
from visitron_b200 import modeling
from visitron_b200.pretrained import BertConfig

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# bf16 activations (eps = 2^-8) with fp32 accumulation against the fp32 reference: stated tolerances
LOGIT_TOL = 4e-2       # absolute, on logits / log-probabilities / hidden states of O(1) magnitude
LOSS_RTOL = 2e-2       # relative, on the scalar losses
GRAD_REL_L2 = 5e-2     # per-tensor relative L2 error of a gradient (sampled or full)
GRAD_COS = 0.998       # per-tensor cosine similarity of a gradient


def make_config(oc: vo.OracleConfig, **kw):
    c = BertConfig(oc.vocab_size, num_hidden_layers=oc.num_hidden_layers, max_position_embeddings=oc.max_position_embeddings,
                   hidden_dropout_prob=oc.hidden_dropout_prob, attention_probs_dropout_prob=oc.attention_probs_dropout_prob)
    c.img_feature_dim = oc.img_feature_dim
    c.use_img_layernorm = oc.use_img_layernorm
    c.detector_classes = oc.detector_classes
    c.action_space = oc.action_space
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def load_params(model, P):
    sd = {k: v.detach().clone() for k, v in P.items()}
    if "mlmhead.predictions.bias" in sd:
        sd["mlmhead.predictions.decoder.weight"] = sd["bert.embeddings.word_embeddings.weight"]
    return model.load_state_dict(sd, strict=True)


def load_fixture(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def digest_close(name, g, dig):
    f = g.detach().float().reshape(-1).cpu()
    if dig["norm"] < 1e-6:  # mathematically zero (key bias: softmax is shift invariant) -- rounding noise on both sides
        assert float(f.norm()) < 1e-3, (name, float(f.norm()))
        return
    assert abs(float(f.norm()) - dig["norm"]) <= GRAD_REL_L2 * dig["norm"], (name, float(f.norm()), dig["norm"])
    s, r = f[::SAMPLE][:512], dig["sample"]
    err = float((s - r).norm() / r.norm().clamp_min(1e-12))
    # a 768-vector is sampled at 3 points only: single elements carry more relative noise than a norm
    assert err < (GRAD_REL_L2 if s.numel() >= 64 else 3 * GRAD_REL_L2), (name, err)


def grad_close(name, mine, ref, floor):
    """Full-tensor gradient comparison: relative L2 and cosine, with an absolute floor for
    gradients that are mathematically (near) zero."""
    a, b = mine.detach().float().reshape(-1), ref.detach().float().reshape(-1)
    nb = float(b.norm())
    if nb < floor:
        assert float(a.norm()) < 20 * floor, (name, float(a.norm()), nb)
        return 0.0, 1.0
    rel = float((a - b).norm()) / nb
    cos = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-30))
    assert rel < GRAD_REL_L2 and cos > GRAD_COS, (name, rel, cos)
    return rel, cos


def decided_argmax_equal(mine, ref_logits, ref_arg, tol=LOGIT_TOL):
    """argmax must agree exactly wherever the fp32 reference's top-2 margin exceeds twice the logit
    tolerance; returns the number of rows that were decided."""
    if ref_logits.size(0) == 0:
        return 0
    top2 = ref_logits.float().topk(2, dim=-1).values
    decided = ((top2[:, 0] - top2[:, 1]) > 2 * tol).to(mine.device)
    assert torch.equal(mine[decided].cpu(), ref_arg.to(mine.device)[decided].cpu())
    return int(decided.sum())


def run_pretrain_fixture(name, device):
    fix = load_fixture(name)
    oc = vo.OracleConfig(**fix["cfg"])
    model = modeling.PreTrainOscar(make_config(oc))
    P = vo.init_params(oc, fix["seed"])
    load_params(model, P)
    model.to(device)
    model.train()
    from visitron_b200 import synthetic
    batch = synthetic.pretrain_batch(**fix["batch_kw"])
    out = model(**{k: v.to(device) for k, v in batch.items()})
    assert len(out) == 7 and all(o.dim() == 0 for o in out)
    out[0].backward()
    for a, b in zip(out, fix["outputs"]):
        assert abs(float(a.detach()) - b) <= LOSS_RTOL * max(1.0, abs(b)), (float(a.detach()), b)
    aux = model._vt_last_aux
    with torch.no_grad():
        _, oaux = vo.pretrain_forward(P, oc, return_aux=True, **batch)
    lab, tl = batch["labels"], batch["token_labels"]
    n = decided_argmax_equal(aux["predicted_words"], oaux["prediction_scores"][lab != -1], fix["predicted_words_labelled"])
    n += decided_argmax_equal(aux["predicted_tokens"], oaux["token_prob"][tl != -1].log(), fix["predicted_tokens_labelled"])
    n += decided_argmax_equal(aux["action_scores"].argmax(1), oaux["action_scores"], fix["action_scores"].argmax(1))
    assert n >= 4, "too many near-ties for the argmax check to mean anything"
    torch.testing.assert_close(aux["mlm_logits"][:, ::16].cpu(), fix["mlm_logits_labelled"], rtol=0, atol=LOGIT_TOL)
    torch.testing.assert_close(aux["action_scores"].cpu(), fix["action_scores"], rtol=0, atol=LOGIT_TOL)
    grads = {k: p.grad for k, p in model.named_parameters()}
    assert set(fix["grads"]) <= set(grads)
    for k, dig in fix["grads"].items():
        assert grads[k] is not None, k
        digest_close(k, grads[k], dig)
    return model


def run_finetune_fixture(device):
    fix = load_fixture("finetune_small")
    oc = vo.OracleConfig(**fix["cfg"])
    model = modeling.BertImgModelwithLocationEmbeds(make_config(oc))
    P = vo.init_params(oc, fix["seed"], pretrain_heads=False)
    model.load_state_dict({k[len("bert."):]: v.clone() for k, v in P.items()}, strict=True)
    model.to(device)
    model.train()
    from visitron_b200 import synthetic
    batch = {k: v.to(device) for k, v in synthetic.finetune_encoder_batch(**fix["batch_kw"]).items()}
    seq, pooled = model(batch["input_ids"], token_type_ids=batch["token_type_ids"], attention_mask=batch["attention_mask"],
                        position_ids=None)[:2]
    assert seq.dtype == torch.float32 and tuple(seq.shape) == tuple(fix["sequence_output"].shape)
    torch.testing.assert_close(seq.detach().cpu(), fix["sequence_output"], rtol=0, atol=6e-2)
    torch.testing.assert_close(pooled.detach().cpu(), fix["pooled_output"], rtol=0, atol=LOGIT_TOL)
    g = torch.Generator().manual_seed(fix["seed"] + 1)
    proj = torch.randn(seq.shape, generator=g) / seq.numel() ** 0.5
    (seq * proj.to(device)).sum().backward()
    grads = {"bert." + k: p.grad for k, p in model.named_parameters()}
    for k, dig in fix["grads"].items():
        digest_close(k, grads[k], dig)
    # parameters this call does not reach keep grad None, as in the reference (find_unused_parameters)
    for k in ("bert.img_embedding.weight", "bert.location_embeds.bias", "bert.pooler.dense.weight"):
        assert grads[k] is None, k
    return model
