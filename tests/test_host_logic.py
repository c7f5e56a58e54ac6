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
import abi_double as emu
from visitron_b200 import modeling, synthetic
from visitron_b200.pretrained import BertConfig

from parity_common import make_config as _cfg, run_finetune_fixture, run_pretrain_fixture


@pytest.mark.parametrize("name", ["pretrain_small", "pretrain_small_imgln"])
def test_pretrain_model_against_reference_fixture(monkeypatch, name):
    emu.install(monkeypatch)
    model = run_pretrain_fixture(name, "cpu")
    # the gradients are views of the flat buffers (no copies were made by autograd)
    st = model.bert._vt_store
    buf = st._grad[st._cur]
    w = model.bert.encoder.layer[1].attention.self.key.weight
    off = st.offset["encoder.layer.1.attention.self.key.weight"]
    assert w.grad.data_ptr() == buf.data_ptr() + 4 * off
    assert "gemm" in emu.CALLS and "attention_bwd" in emu.CALLS


def test_finetune_encoder_against_reference_fixture(monkeypatch):
    emu.install(monkeypatch)
    run_finetune_fixture("cpu")


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


def test_parameter_used_later_keeps_accumulated_gradient(monkeypatch):
    """a parameter that gets its FIRST gradient in a later step (the next-action head when `next_action` is None in
    step 1) with in-place zeroing ends up in another gradient buffer than its siblings; no later backward may wipe a live .grad, and
    gradient accumulation must add, not double."""
    emu.install(monkeypatch)
    oc = vo.OracleConfig(vocab_size=300, num_hidden_layers=1, max_position_embeddings=16, detector_classes=20)
    model = modeling.PreTrainOscar(_cfg(oc))
    model.train()
    batch = synthetic.pretrain_batch(2, 3, text_len=12, regions=8, vocab=300, detector_classes=20)
    no_action = {k: v for k, v in batch.items() if k != "next_action"}

    def grads():
        return {k: (None if p.grad is None else p.grad.clone()) for k, p in model.named_parameters()}

    model(**batch)[0].backward()
    ref = grads()                                       # one full step from scratch
    for p in model.parameters():
        p.grad = None
    model(**no_action)[0].backward()                    # step 1: the action head (and the pooler) get no gradient
    assert model.next_action.linear.weight.grad is None
    g_no = grads()
    for p in model.parameters():                        # zero_grad(set_to_none=False)
        if p.grad is not None:
            p.grad.zero_()
    model(**batch)[0].backward()                        # step 2: first gradient of the action head, the others accumulate into zeros
    model(**batch)[0].backward()                        # step 3: accumulation on top of step 2
    for k, p in model.named_parameters():
        if ref[k] is None:
            continue
        torch.testing.assert_close(p.grad, 2 * ref[k], rtol=1e-4, atol=1e-6, msg=k)
    assert g_no["bert.encoder.layer.0.output.dense.weight"] is not None


def test_frozen_encoder_with_trainable_heads(monkeypatch):
    """the heads must save their activations when only THEY require gradients."""
    emu.install(monkeypatch)
    oc = vo.OracleConfig(vocab_size=300, num_hidden_layers=1, max_position_embeddings=16, detector_classes=20)
    model = modeling.PreTrainOscar(_cfg(oc))
    model.train()
    for p in model.bert.parameters():
        p.requires_grad_(False)
    batch = synthetic.pretrain_batch(2, 3, text_len=12, regions=8, vocab=300, detector_classes=20)
    model(**batch)[0].backward()
    assert model.token_head[0].weight.grad is not None and model.mlmhead.predictions.transform.dense.weight.grad is not None
    assert model.bert.encoder.layer[0].output.dense.weight.grad is None


def test_exchange_cta_cap_follows_the_world_size(monkeypatch):
    """parallel._max_ctas: 4 NCCL CTAs up to 3 ranks, 8 from 4 ranks on (measured, see parallel.py); VT_NCCL_MAX_CTAS overrides."""
    from visitron_b200 import parallel
    monkeypatch.setattr(parallel, "NCCL_MAX_CTAS", 0)
    assert [parallel._max_ctas(w) for w in (1, 2, 3, 4, 8)] == [4, 4, 4, 8, 8]
    monkeypatch.setattr(parallel, "NCCL_MAX_CTAS", 6)
    assert parallel._max_ctas(2) == 6 and parallel._max_ctas(8) == 6
