"""Sync-free step machinery on the B200: device-side label compaction, device-resident row limits, device-resident
dropout seeds, the captured CUDA-graph step, and the bf16 operand refresh (stale operands under the
reference's AdamW).  Everything goes through the C ABI."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from visitron_b200 import flat, graph as vgraph, lib, modeling, ops, optimization as vopt, synthetic

from oracle import visitron_oracle as vo
from parity_common import load_params, make_config

DEV = "cuda"
BF = torch.bfloat16


def _small(dropout=0.0, layers=2):
    oc = vo.OracleConfig(vocab_size=2048, num_hidden_layers=layers, max_position_embeddings=64, detector_classes=96,
                         hidden_dropout_prob=dropout, attention_probs_dropout_prob=dropout / 3)
    model = modeling.PreTrainOscar(make_config(oc))
    load_params(model, vo.init_params(oc, 5))
    return oc, model.to(DEV).train()


def _batch(b=3, seed=88):
    return {k: v.to(DEV) for k, v in synthetic.pretrain_batch(b, seed, text_len=40, regions=24, vocab=2048, detector_classes=96).items()}


@pytest.mark.parametrize("n,frac", [(24544, 0.06), (1000, 0.5), (130, 1.0), (64, 0.0), (1, 1.0)])
def test_compact_labels_matches_nonzero(n, frac):
    g = torch.Generator().manual_seed(n)
    a = torch.where(torch.rand(n, generator=g) < frac, torch.randint(0, 2048, (n,), generator=g), torch.tensor(-1))
    b = torch.where(torch.rand(n, generator=g) < frac / 2, torch.randint(0, 96, (n,), generator=g), torch.tensor(-1))
    cap = ops.roundup(n, 128)
    ia, la, ib, lb, counts = ops.compact_labels(a.to(DEV), 2048, b.to(DEV), 96, cap)
    c = counts.cpu().tolist()
    for lab, idx, val, off in ((a, ia, la, 0), (b, ib, lb, 4)):
        ref = torch.nonzero(lab != -1).squeeze(1)
        k = ref.numel()
        assert c[off + 1] == k and c[off] == min(cap, (k + 127) // 128 * 128) and c[off + 2] == 0
        assert torch.equal(idx[:k].cpu(), ref) and torch.equal(val[:k].cpu(), lab[ref])
        assert bool((idx[k:c[off]] == -1).all()) and bool((val[k:c[off]] == 0).all())


def test_compact_labels_reports_out_of_range_and_loss_turns_nan():
    oc, model = _small()
    batch = _batch()
    batch["labels"] = batch["labels"].clone()
    pos = int(torch.nonzero(batch["labels"].reshape(-1) != -1)[0])
    batch["labels"].view(-1)[pos] = 2048 + 7          # torch's CrossEntropyLoss raises a device assert here
    out = model(**batch)
    assert torch.isnan(out[0]) and model._vt_last_aux["label_errors"] == 1


def test_row_limit_clips_gemm_and_row_kernels():
    torch.manual_seed(3)
    cap, n, N, K = 512, 200, 320, 768
    counts = torch.tensor([256, n, 0, 0, 0, 0, 0, 0], dtype=torch.int32, device=DEV)
    A = (torch.randn(cap, K, device=DEV) * 0.3).to(BF)
    A[n:256] = 0                                       # the padding rows are zeros (gather of idx = -1)
    W = (torch.randn(N, K, device=DEV) * 0.05).to(BF)
    out = torch.full((cap, N), 7.0, device=DEV, dtype=BF)
    with ops.row_limit(counts, 0):
        lib.gemm(A, W, out=out)
    ref = (A[:256].float() @ W.float().t())
    assert float((out[:256].float() - ref).abs().max()) < 2e-2 * float(ref.abs().max())
    assert bool((out[256:] == 7.0).all()), "rows past the padded count must not be touched"
    # weight-gradient form: the contraction runs over the limited rows; stale rows past the padded count must not count
    dY = (torch.randn(cap, N, device=DEV) * 0.3).to(BF)
    dY[n:256] = 0
    A2 = A.clone()
    A2[256:] = 1000.0
    dW = torch.zeros(N, K, device=DEV)
    with ops.row_limit(counts, 0):
        lib.gemm(dY, A2, a_major=1, b_major=1, out=dW, accumulate=True)
    refw = dY[:n].float().t() @ A[:n].float()
    assert float((dW - refw).norm() / refw.norm()) < 1e-2
    # an empty count: nothing runs, nothing is written
    zero = torch.zeros(8, dtype=torch.int32, device=DEV)
    dW2 = torch.zeros(N, K, device=DEV)
    out2 = torch.full((cap, N), 7.0, device=DEV, dtype=BF)
    with ops.row_limit(zero, 0):
        lib.gemm(dY, A2, a_major=1, b_major=1, out=dW2, accumulate=True)
        lib.gemm(A, W, out=out2)
        cs = ops.colsum(dY, torch.zeros(N, device=DEV))
    assert float(dW2.abs().max()) == 0.0 and bool((out2 == 7.0).all()) and float(cs.abs().max()) == 0.0
    # column sums and LayerNorm under the limit
    with ops.row_limit(counts, 0):
        cs = ops.colsum(dY, torch.zeros(N, device=DEV))
        g, bta = torch.ones(768, device=DEV), torch.zeros(768, device=DEV)
        y = ops.layernorm_fwd(A, g, bta, 1e-12, out=torch.full((cap, K), 3.0, device=DEV, dtype=BF))
    assert float((cs - dY[:256].float().sum(0)).abs().max()) < 1e-2
    assert bool((y[256:] == 3.0).all()) and bool(torch.isfinite(y[:256].float()).all())


def test_step_has_no_host_sync():
    """The forward + backward of PreTrainOscar enqueues without waiting for the device: under
    torch.cuda.set_sync_debug_mode("error") any synchronising call raises."""
    oc, model = _small(dropout=0.1)
    batch = _batch()
    out = model(**batch)       # first call: lazy initialisation may sync
    out[0].backward()
    torch.cuda.synchronize()
    for p in model.parameters():
        p.grad = None
    torch.cuda.set_sync_debug_mode("error")
    try:
        out = model(**batch)
        out[0].backward()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    torch.cuda.synchronize()
    assert torch.isfinite(out[0])


def test_graphed_step_equals_eager_step():
    oc, model = _small(dropout=0.0)
    batches = [_batch(3, 88 + i) for i in range(3)]
    eager = []
    for bt in batches:
        for p in model.parameters():
            p.grad = None
        out = model(**bt)
        out[0].backward()
        eager.append((torch.stack([o.detach() for o in out]).cpu(), {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}))
    del out
    step = vgraph.GraphedStep(model, batches[0])
    assert step.launches_per_step > 50
    for bt, (ref_out, ref_g) in zip(batches + batches[:1], eager + eager[:1]):   # different label counts per batch, and a re-run
        for p in model.parameters():
            p.grad = None                                # the loop's zero_grad(set_to_none=True)
        out = step(bt)
        got = torch.stack(list(out)).cpu()
        torch.testing.assert_close(got, ref_out, rtol=1e-5, atol=1e-6, equal_nan=True)
        for k, p in model.named_parameters():
            if k in ref_g:
                assert p.grad is not None, k
                d = float((p.grad - ref_g[k]).norm()) / max(float(ref_g[k].norm()), 1e-12)
                assert d < 1e-4, (k, d)                  # fp32 red.add order of the split-K weight gradients


def test_graphed_step_draws_fresh_dropout_masks_and_is_reproducible():
    oc, model = _small(dropout=0.3)
    bt = _batch()
    s1 = vgraph.GraphedStep(model, bt, seed=1234)
    a = [float(s1(bt)[0]) for _ in range(4)]
    assert len(set(a)) == 4, a                           # each replay reads a new seed from device memory
    s2 = vgraph.GraphedStep(model, bt, seed=1234)
    b = [float(s2(bt)[0]) for _ in range(4)]
    assert a == b, (a, b)                                # same seed, same mask sequence
    s3 = vgraph.GraphedStep(model, bt, seed=99)
    assert float(s3(bt)[0]) != a[0]
    # eager calls beside the graph keep their by-value seeds
    torch.manual_seed(5)
    e1 = float(model(**bt)[0])
    torch.manual_seed(5)
    e2 = float(model(**bt)[0])
    assert e1 == e2


def test_weights_changed_through_p_data_reach_the_gemms():
    """the reference's AdamW updates with p.data.add_ / addcdiv_
    (transformers/pytorch_transformers/optimization.py:176,187), which does not move the autograd version counters."""
    oc, model = _small()
    bt = _batch()
    l0 = float(model(**bt)[0])
    with torch.no_grad():
        for k, p in model.named_parameters():
            if "intermediate.dense.weight" in k or "word_embeddings" in k:
                v = p._version
                p.data.add_(0.05 * torch.randn_like(p))
                assert p._version == v
    l1 = float(model(**bt)[0])
    assert abs(l1 - l0) > 1e-3, (l0, l1)
    fresh = modeling.PreTrainOscar(model.config)
    fresh.load_state_dict(model.state_dict())
    l2 = float(fresh.to(DEV).train()(**bt)[0])
    assert abs(l1 - l2) < 1e-5, (l1, l2)


def test_fused_adamw_writes_the_bf16_operands():
    oc, model = _small()
    bt = _batch()
    opt = vopt.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
    for it in range(2):
        opt.zero_grad(set_to_none=True)
        model(**bt)[0].backward()
        opt.step()
        for st in (model.bert._vt_store, model._vt_store):
            assert st._fresh is not None
            assert torch.equal(st.bf16.view(torch.int16), st.flat.to(BF).view(torch.int16)), "bf16 shadow != cast(fp32 master) after the fused step"
    n0 = lib.launch_count()
    model(**bt)
    with_note = lib.launch_count() - n0
    n0 = lib.launch_count()
    model(**bt)                                          # the note is single-use: this forward casts again
    without = lib.launch_count() - n0
    assert without == with_note + 2, (with_note, without)
    flat.invalidate()


def test_untied_decoder_after_resize_embeddings_trains():
    """vs/model_utils.py:101-109 (--no_oscar_setting): resize_embeddings un-ties the MLM decoder (vs/encoder.py:337-345);
    the decoder then is a parameter of its own with its own gradient."""
    oc, model = _small()
    bt = _batch()
    ref = vo.init_params(oc, 5)
    model.resize_embeddings({"word_embeddings": 2048 + 3, "token_type_embeddings": 2 + 4})
    model.to(DEV)
    dec = model.mlmhead.predictions.decoder.weight
    assert dec is not model.bert.embeddings.word_embeddings.weight and dec.size(0) == 2048
    out = model(**bt)
    out[0].backward()
    P = {k: v.clone() for k, v in ref.items()}
    sd = model.state_dict()
    for k in ("bert.embeddings.word_embeddings.weight", "bert.embeddings.token_type_embeddings.weight"):
        P[k] = sd[k].detach().cpu().clone()
    P["mlmhead.predictions.decoder.weight"] = sd["mlmhead.predictions.decoder.weight"].detach().cpu().clone()
    Pg = vo.clone_params(P)
    r = vo.pretrain_forward(Pg, oc, **{k: v.cpu() for k, v in bt.items()})
    r[0].backward()
    for a, b in zip(out, r):
        assert abs(float(a) - float(b)) <= 2e-2 * max(1.0, abs(float(b)))
    for k in ("mlmhead.predictions.decoder.weight", "bert.embeddings.word_embeddings.weight"):
        g, gr = dict(model.named_parameters())[k].grad.cpu(), Pg[k].grad
        assert float((g - gr).norm() / gr.norm()) < 5e-2, k


def test_output_hidden_states():
    oc = vo.OracleConfig(vocab_size=2048, num_hidden_layers=2, max_position_embeddings=64, detector_classes=96)
    cfg = make_config(oc)
    cfg.output_hidden_states = True
    model = modeling.BertImgModelwithLocationEmbeds(cfg).to(DEV).eval()
    ids = torch.randint(1, 2048, (2, 40), device=DEV)
    out = model(ids)
    assert len(out) == 3 and len(out[2]) == 3 and all(h.shape == (2, 40, 768) for h in out[2])
    assert torch.equal(out[2][-1], out[0].detach())
