"""Host-side logic of visitron_b200/modeling.py on CPU: the model classes are driven through the
TEST DOUBLE of the C ABI (tests/abi_double.py, plain torch restatements of every wrapper) and compared with
the oracle / the reference's golden fixtures.  This checks what Python decides -- operand majors,
flat-buffer offsets, residual and remap wiring, gradient bookkeeping, API surface -- not the CUDA
kernels (those are the ``-m gpu`` tests)."""
import copy
import os

import pytest
import torch

from oracle import visitron_oracle as vo
from oracle.make_golden import SAMPLE
import abi_double as emu
from visitron_b200 import modeling, synthetic
from visitron_b200.pretrained import BertConfig

LOGIT_TOL = 4e-2  # bf16 activations (eps 2^-8) through 2 layers + head vs the fp32 reference; stated tolerance for logits
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _cfg(oc: vo.OracleConfig, **kw):
    c = BertConfig(oc.vocab_size, num_hidden_layers=oc.num_hidden_layers, max_position_embeddings=oc.max_position_embeddings,
                   hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    c.img_feature_dim = oc.img_feature_dim
    c.use_img_layernorm = oc.use_img_layernorm
    c.detector_classes = oc.detector_classes
    c.action_space = oc.action_space
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _load_params(model, P):
    sd = {k: v.clone() for k, v in P.items()}
    if "mlmhead.predictions.bias" in sd:
        sd["mlmhead.predictions.decoder.weight"] = sd["bert.embeddings.word_embeddings.weight"]
    missing = model.load_state_dict(sd, strict=True)
    return missing


def _digest_close(name, g, dig, rel=3e-2):
    f = g.detach().float().reshape(-1)
    if dig["norm"] < 1e-6:  # mathematically zero (key bias: softmax is shift invariant) -- only rounding noise on both sides
        assert float(f.norm()) < 1e-4, (name, float(f.norm()))
        return
    assert abs(float(f.norm()) - dig["norm"]) <= rel * max(dig["norm"], 1e-6), (name, float(f.norm()), dig["norm"])
    s, r = f[::SAMPLE][:512], dig["sample"]
    err = float((s - r).norm() / r.norm().clamp_min(1e-12))
    assert err < 5e-2, (name, err)


@pytest.mark.parametrize("name", ["pretrain_small", "pretrain_small_imgln"])
def test_pretrain_model_against_reference_fixture(monkeypatch, name):
    emu.install(monkeypatch)
    fix = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    oc = vo.OracleConfig(**fix["cfg"])
    model = modeling.PreTrainOscar(_cfg(oc))
    _load_params(model, vo.init_params(oc, fix["seed"]))
    model.train()
    batch = synthetic.pretrain_batch(**fix["batch_kw"])
    out = model(**batch)
    assert len(out) == 7 and all(o.dim() == 0 for o in out)
    out[0].backward()
    for a, b in zip(out, fix["outputs"]):
        assert abs(float(a.detach()) - b) <= 2e-2 * max(1.0, abs(b)), (float(a.detach()), b)
    aux = model._vt_last_aux
    # argmax must agree exactly wherever the fp32 oracle's top-2 margin exceeds the bf16 logit tolerance
    P = vo.init_params(oc, fix["seed"])
    with torch.no_grad():
        _, oaux = vo.pretrain_forward(P, oc, return_aux=True, **batch)
    lab, tl = batch["labels"], batch["token_labels"]
    n_decided = 0
    for mine, ref_logits, ref_arg, tol in (
            (aux["predicted_words"], oaux["prediction_scores"][lab != -1], fix["predicted_words_labelled"], LOGIT_TOL),
            (aux["predicted_tokens"], oaux["token_prob"][tl != -1].log(), fix["predicted_tokens_labelled"], LOGIT_TOL),
            (aux["action_scores"].argmax(1), oaux["action_scores"], fix["action_scores"].argmax(1), LOGIT_TOL)):
        top2 = ref_logits.topk(2, dim=-1).values
        decided = (top2[:, 0] - top2[:, 1]) > 2 * tol
        n_decided += int(decided.sum())
        assert torch.equal(mine[decided], ref_arg[decided])
    assert n_decided >= 4, "too many near-ties for the check to mean anything"
    torch.testing.assert_close(aux["mlm_logits"][:, ::16], fix["mlm_logits_labelled"], rtol=0, atol=LOGIT_TOL)
    torch.testing.assert_close(aux["action_scores"], fix["action_scores"], rtol=0, atol=3e-2)
    grads = {k: p.grad for k, p in model.named_parameters()}
    assert set(fix["grads"]) <= set(grads)
    for k, dig in fix["grads"].items():
        assert grads[k] is not None, k
        _digest_close(k, grads[k], dig)
    # the gradients are views of the flat buffers (no copies were made by autograd)
    st = model.bert._vt_store
    buf = st._grad[st._cur]
    w = model.bert.encoder.layer[1].attention.self.key.weight
    off = st.offset["encoder.layer.1.attention.self.key.weight"]
    assert w.grad.data_ptr() == buf.data_ptr() + 4 * off
    assert "gemm" in emu.CALLS and "attention_bwd" in emu.CALLS


def test_finetune_encoder_against_reference_fixture(monkeypatch):
    emu.install(monkeypatch)
    fix = torch.load(os.path.join(GOLDEN, "finetune_small.pt"), weights_only=False)
    oc = vo.OracleConfig(**fix["cfg"])
    model = modeling.BertImgModelwithLocationEmbeds(_cfg(oc))
    P = vo.init_params(oc, fix["seed"], pretrain_heads=False)
    model.load_state_dict({k[len("bert."):]: v.clone() for k, v in P.items()}, strict=True)
    model.train()
    batch = synthetic.finetune_encoder_batch(**fix["batch_kw"])
    seq, pooled = model(batch["input_ids"], token_type_ids=batch["token_type_ids"], attention_mask=batch["attention_mask"],
                        position_ids=None)[:2]
    assert seq.dtype == torch.float32 and tuple(seq.shape) == tuple(fix["sequence_output"].shape)
    torch.testing.assert_close(seq, fix["sequence_output"], rtol=0, atol=6e-2)
    torch.testing.assert_close(pooled, fix["pooled_output"], rtol=0, atol=3e-2)
    g = torch.Generator().manual_seed(fix["seed"] + 1)
    proj = torch.randn(seq.shape, generator=g) / seq.numel() ** 0.5
    (seq * proj).sum().backward()
    grads = {"bert." + k: p.grad for k, p in model.named_parameters()}
    for k, dig in fix["grads"].items():
        _digest_close(k, grads[k], dig)
    # parameters this call does not reach keep grad None, as in the reference (find_unused_parameters)
    for k in ("bert.img_embedding.weight", "bert.location_embeds.bias", "bert.pooler.dense.weight"):
        assert grads[k] is None, k


def test_state_dict_keys_match_the_reference_layout():
    oc = vo.OracleConfig(vocab_size=512, num_hidden_layers=2, max_position_embeddings=32, detector_classes=40, use_img_layernorm=1)
    model = modeling.PreTrainOscar(_cfg(oc))
    want = set(vo.param_shapes(oc)) | {"mlmhead.predictions.decoder.weight"}
    sd = model.state_dict()
    assert set(sd) == want
    for k, shp in vo.param_shapes(oc).items():
        assert tuple(sd[k].shape) == shp, k
    assert model.mlmhead.predictions.decoder.weight is model.bert.embeddings.word_embeddings.weight
    # optimizer grouping by name (vs/pretrain.py:109-127)
    no_decay = [n for n, _ in model.named_parameters() if any(nd in n for nd in ("bias", "LayerNorm.weight"))]
    assert "token_head.0.bias" in no_decay and "bert.LayerNorm.weight" in no_decay and "mlmhead.predictions.bias" in no_decay


def test_flat_store_survives_optimizer_steps_and_moves(monkeypatch):
    emu.install(monkeypatch)
    oc = vo.OracleConfig(vocab_size=300, num_hidden_layers=1, max_position_embeddings=16, detector_classes=20)
    model = modeling.PreTrainOscar(_cfg(oc))
    model.train()
    batch = synthetic.pretrain_batch(2, 3, text_len=12, regions=8, vocab=300, detector_classes=20)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    losses = []
    for it in range(3):
        opt.zero_grad()
        out = model(**batch)
        out[0].backward()
        opt.step()
        losses.append(float(out[0]))
    assert losses[2] < losses[0]                       # the bf16 operand cache follows in-place parameter updates
    st = model.bert._vt_store
    assert st._aliased()
    sd = copy.deepcopy(model.state_dict())
    m2 = modeling.PreTrainOscar(_cfg(oc))
    m2.load_state_dict(sd)
    m2.eval()
    model.eval()
    with torch.no_grad():
        a, b = model(**batch), m2(**batch)
    assert abs(float(a[0]) - float(b[0])) < 1e-6
    # zero_grad(set_to_none=False): the old .grad still aliases buffer 0, so the next backward must use buffer 1
    model.train()
    opt.zero_grad(set_to_none=False)
    out = model(**batch)
    out[0].backward()
    g1 = model.bert.pooler.dense.weight.grad.clone()
    opt.zero_grad(set_to_none=True)
    out = model(**batch)
    out[0].backward()
    torch.testing.assert_close(model.bert.pooler.dense.weight.grad, g1, rtol=1e-5, atol=1e-7)


def test_unsupported_options_raise(monkeypatch):
    emu.install(monkeypatch)
    oc = vo.OracleConfig(vocab_size=300, num_hidden_layers=1, max_position_embeddings=16, detector_classes=20)
    model = modeling.BertImgModelwithLocationEmbeds(_cfg(oc))
    ids = torch.randint(1, 300, (2, 8))
    with pytest.raises(NotImplementedError):
        model(ids, head_mask=torch.ones(12))
    with pytest.raises(NotImplementedError):
        model(ids, attention_mask=torch.ones(2, 8, 8))
    with pytest.raises(NotImplementedError):
        model(ids, attention_mask=torch.ones(2, 1, 1, 8))
    with pytest.raises(NotImplementedError):
        modeling.BertImgModelwithLocationEmbeds(_cfg(oc, hidden_size=512, num_attention_heads=8))
    with pytest.raises(ValueError):
        modeling.BertImgModelwithLocationEmbeds(_cfg(oc, num_attention_heads=7))


def test_product_refuses_cpu_tensors_without_the_double():
    from visitron_b200 import lib
    oc = vo.OracleConfig(vocab_size=300, num_hidden_layers=1, max_position_embeddings=16, detector_classes=20)
    model = modeling.BertImgModelwithLocationEmbeds(_cfg(oc))
    with pytest.raises(lib.VtError):
        model(torch.randint(1, 300, (2, 8)))
