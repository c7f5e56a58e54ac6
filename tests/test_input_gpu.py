"""The step before forward on the B200 kernel, through the C ABI: masked-LM masking of a whole batch in one
launch against the oracle restatement of PretrainDataset._mask_tokens (itself pinned to the reference by
tests/golden/mlm_mask_small.pt) run on the kernel's own draws.  Integer work: bit-exact."""
import pytest
import torch

from oracle import visitron_oracle as vo
from parity_common import load_fixture
from visitron_b200 import input_pipeline as ip, lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
V, MASK, PAD, SPECIAL = 30522, 103, 0, [100, 102, 0, 101, 103]


def _batch(b, L, seed, with_classes=True):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1000, V, (b, L), generator=g)
    tc = torch.full((b, L), -1, dtype=torch.long)
    ids[:, 0] = 101
    for r in range(b):
        n = int(torch.randint(L // 4, L + 1, (1,), generator=g)) if L > 4 else L
        ids[r, n - 1] = 102
        ids[r, n:] = PAD
        if with_classes and n > 4:
            pos = torch.randperm(n - 2, generator=g)[: max(1, n // 16)] + 1
            tc[r, pos] = torch.randint(0, 1601, (pos.numel(),), generator=g)
    return ids, (tc if with_classes else None)


def _oracle(ids, tc, seed, p=0.15, vocab=V, special=SPECIAL):
    u, r = ip.draws(ids.numel(), vocab, seed, DEV)
    u, r = u.cpu().reshape(3, *ids.shape), r.cpu().reshape(ids.shape)
    sp = torch.zeros(vocab, dtype=torch.bool)
    sp[torch.tensor(special)] = True
    return vo.mlm_mask(ids, tc, sp[ids], u[0] < p, u[1] < 0.8, u[2] < 0.5, r, MASK, PAD), u, r


@pytest.mark.parametrize("b,L,classes", [(32, 511, True), (3, 80, False), (1, 1, True), (7, 333, True)])
def test_mlm_mask_matches_oracle_bit_exact(b, L, classes):
    ids, tc = _batch(b, L, 11 + b, classes)
    lut = ip.special_lut(V, SPECIAL, DEV)
    n0 = lib.launch_count()
    out, labels, att = ip.mask_tokens(ids.to(DEV), None if tc is None else tc.to(DEV), lut, MASK, PAD, seed=1234 + L)
    assert lib.launch_count() - n0 == 1, "the whole batch in ONE launch"
    (w_ids, w_labels, w_att), _, _ = _oracle(ids, tc, 1234 + L)
    assert torch.equal(out.cpu(), w_ids) and torch.equal(labels.cpu(), w_labels) and torch.equal(att.cpu(), w_att)
    assert att.dtype == torch.bool and out.dtype == torch.int64 and labels.dtype == torch.int64


def test_mlm_mask_on_the_reference_fixture_rows():
    """The rows (and vocabulary) of the reference fixture through the kernel: same rule as the reference's outputs wherever the draws
    agree is covered by the oracle test; here the kernel == oracle on those very rows, and the rule's invariants hold."""
    fix = load_fixture("mlm_mask_small")
    lut = ip.special_lut(fix["vocab"], fix["special"], DEV)
    for k, c in enumerate(fix["cases"]):
        tc = c["token_classes"] if c["mtp"] else None
        out, labels, att = ip.mask_tokens(c["ids"].to(DEV), None if tc is None else tc.to(DEV), lut, fix["mask_id"], fix["pad_id"], seed=k)
        (w_ids, w_labels, w_att), _, _ = _oracle(c["ids"], tc, k, vocab=fix["vocab"], special=fix["special"])
        assert torch.equal(out.cpu(), w_ids) and torch.equal(labels.cpu(), w_labels) and torch.equal(att.cpu(), w_att)
        assert torch.equal(att.cpu(), c["attention"]), "attention mask does not depend on the draws"
        sp = torch.isin(c["ids"], torch.tensor(fix["special"]))
        free = sp & (c["token_classes"] == -1) if c["mtp"] else sp
        assert bool((labels.cpu()[sp] == -1).all()) and torch.equal(out.cpu()[free], c["ids"][free])


def test_mlm_mask_statistics_and_determinism_full_batch():
    """BASELINE-size text block (b=256 x 511 tokens): 15 % chosen, 80 / 10 / 10 split, specials untouched, forced region-label
    positions, same seed -> same result, another seed -> another mask; the input tensor is left alone."""
    ids, tc = _batch(256, 511, 5, True)
    d_ids, d_tc = ids.to(DEV), tc.to(DEV)
    lut = ip.special_lut(V, SPECIAL, DEV)
    keep = d_ids.clone()
    out, labels, att = ip.mask_tokens(d_ids, d_tc, lut, MASK, PAD, seed=99)
    out2, labels2, _ = ip.mask_tokens(d_ids, d_tc, lut, MASK, PAD, seed=99)
    out3, _, _ = ip.mask_tokens(d_ids, d_tc, lut, MASK, PAD, seed=100)
    assert torch.equal(d_ids, keep)
    assert torch.equal(out, out2) and torch.equal(labels, labels2) and not torch.equal(out, out3)
    out, labels = out.cpu(), labels.cpu()
    special = torch.isin(ids, torch.tensor(SPECIAL))
    forced = tc != -1
    assert bool((out[forced] == MASK).all()) and bool((labels[forced] == -1).all())
    assert torch.equal(out[special & ~forced], ids[special & ~forced]) and bool((labels[special] == -1).all())
    assert torch.equal(att.cpu(), ids != PAD)
    cand = ~special & ~forced
    chosen = labels != -1
    assert bool((labels[chosen] == ids[chosen]).all()) and not bool((chosen & ~cand).any())
    rate = chosen.sum().item() / cand.sum().item()
    assert abs(rate - 0.15) < 0.005, rate
    n = chosen.sum().item()
    as_mask = (out[chosen] == MASK).sum().item() / n
    changed = ((out[chosen] != MASK) & (out[chosen] != ids[chosen])).sum().item() / n
    assert abs(as_mask - 0.8) < 0.01 and abs(changed - 0.1) < 0.01, (as_mask, changed)
    assert bool((out[~chosen & ~forced] == ids[~chosen & ~forced]).all())
    rnd = out[chosen & (out != MASK) & (out != ids)]
    assert int(rnd.min()) >= 0 and int(rnd.max()) < V and rnd.float().mean().item() == pytest.approx(V / 2, rel=0.05)


def test_mlm_mask_refuses_cpu_and_bad_dtypes():
    lut = ip.special_lut(V, SPECIAL, DEV)
    with pytest.raises(lib.VtError):
        ip.mask_tokens(torch.zeros(4, dtype=torch.long), None, lut, MASK, PAD, seed=0)
    with pytest.raises(lib.VtError):
        ip.mask_tokens(torch.zeros(4, dtype=torch.int32, device=DEV), None, lut, MASK, PAD, seed=0)
    out, labels, att = ip.mask_tokens(torch.zeros((0, 8), dtype=torch.long, device=DEV), None, lut, MASK, PAD, seed=0)
    assert out.shape == (0, 8) and labels.shape == (0, 8) and att.shape == (0, 8)


# ---- region packing and the whole-batch collation -----------------------------------------------------------------------
def _ragged(counts, D, DL, seed):
    g = torch.Generator().manual_seed(seed)
    n = int(sum(counts))
    feats, loc = torch.randn(n, D, generator=g), torch.randn(n, DL, generator=g)
    off = torch.zeros(len(counts) + 1, dtype=torch.long)
    off[1:] = torch.tensor(counts, dtype=torch.long).cumsum(0)
    return feats, loc, off


@pytest.mark.parametrize("counts,R,D,DL", [([180, 256, 300, 0, 1, 255, 257, 90], 256, 2054, 128), ([3, 0, 9], 4, 7, 5), ([0, 0], 3, 8, 2),
                                           ([5], 1, 2054, 2052)])
def test_pack_regions_matches_oracle_bit_exact(counts, R, D, DL):
    feats, loc, off = _ragged(counts, D, DL, 3)
    B, T = len(counts), 5
    att = torch.full((B, T + R), 7, dtype=torch.long, device=DEV)
    n0 = lib.launch_count()
    img, lo = ip.pack_regions(feats.to(DEV), loc.to(DEV), off.to(DEV), R, attention_out=att, attention_col=T)
    assert lib.launch_count() - n0 == 1
    dummy = (torch.zeros(T, dtype=torch.long), torch.zeros(T, dtype=torch.long), torch.full((T,), 7))
    for b, n in enumerate(counts):
        a, e = int(off[b]), int(off[b + 1])
        want = vo.preprocess_item(dummy, torch.zeros(T, dtype=torch.long), feats[a:e], loc[a:e], 0, R)
        assert torch.equal(img[b].cpu(), want["img_feats"]) and torch.equal(lo[b].cpu(), want["img_location_embeddings"]), b
        assert att[b].tolist() == want["attention_mask"], b


def test_collate_pretrain_on_the_reference_fixture():
    """The samples of the reference's _preprocess_item fixture as ONE device batch: features, locations, attention and token labels equal
    the reference's outputs bit for bit (they do not depend on the random draws); ids / labels equal the oracle on the kernel's draws."""
    fix = load_fixture("preprocess_item_small")
    R = fix["R"]
    for mtp in (True, False):
        cs = [c for c in fix["cases"] if c["mtp"] == mtp and not c["no_action_grounding"]]
        ids = torch.stack([c["ids"] for c in cs])
        tc = torch.stack([c["token_classes"] for c in cs])
        feats, loc = torch.cat([c["feats"] for c in cs]), torch.cat([c["loc"] for c in cs])
        off = torch.tensor([0] + [c["n"] for c in cs]).cumsum(0)
        lut = ip.special_lut(fix["vocab"], fix["special"], DEV)
        na = torch.tensor([c["next_action"] for c in cs])
        batch = ip.collate_pretrain(ids.to(DEV), tc.to(DEV), feats.to(DEV), loc.to(DEV), off.to(DEV), na.to(DEV), lut, fix["mask_id"], fix["pad_id"],
                                    seed=5, max_img_seq_length=R, masked_token_prediction=mtp)
        (w_ids, w_labels, _), _, _ = _oracle(ids, tc if mtp else None, 5, vocab=fix["vocab"], special=fix["special"])
        T = ids.shape[1]
        assert torch.equal(batch["input_ids"].cpu(), w_ids) and torch.equal(batch["labels"][:, :T].cpu(), w_labels)
        assert bool((batch["labels"][:, T:] == -1).all())
        for b, c in enumerate(cs):
            want = c["out"]
            assert torch.equal(batch["img_feats"][b].cpu(), want["img_feats"])
            assert torch.equal(batch["img_location_embeddings"][b].cpu(), want["img_location_embeddings"])
            assert batch["attention_mask"][b].tolist() == [int(a) for a in want["attention_mask"]]
            assert int(batch["next_action"][b]) == want["next_action"]
            if mtp:
                assert torch.equal(batch["token_labels"][b].cpu(), want["token_labels"])
            else:
                assert batch["token_labels"] is None and want["token_labels"] is None
    nag = ip.collate_pretrain(ids.to(DEV), tc.to(DEV), feats.to(DEV), loc.to(DEV), off.to(DEV), na.to(DEV), lut, fix["mask_id"], fix["pad_id"], seed=5,
                              max_img_seq_length=R, no_action_grounding=True)
    assert bool((nag["next_action"] == -1).all())


def test_collated_batch_runs_through_the_model():
    """collate_pretrain's dictionary is what PreTrainOscar.forward takes (same keys as the reference's loader output)."""
    from visitron_b200 import modeling
    from visitron_b200.pretrained import BertConfig
    cfg = BertConfig(30522, num_hidden_layers=1, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    cfg.img_feature_dim, cfg.detector_classes, cfg.action_space = 2054, 1601, 36
    torch.manual_seed(0)
    model = modeling.PreTrainOscar(cfg).to(DEV).train()
    ids, tc = _batch(4, 64, 21, True)
    feats, loc, off = _ragged([10, 32, 40, 0], 2054, 128, 4)
    lut = ip.special_lut(V, SPECIAL, DEV)
    batch = ip.collate_pretrain(ids.to(DEV), tc.to(DEV), feats.to(DEV), loc.to(DEV), off.to(DEV), torch.tensor([1, 5, 35, 0], device=DEV), lut, MASK,
                                PAD, seed=8, max_img_seq_length=32)
    out = model(**batch)
    out[0].backward()
    assert torch.isfinite(out[0]) and model.bert.embeddings.word_embeddings.weight.grad is not None


def test_device_prefetcher_yields_every_batch_in_order():
    host = [dict(a=torch.full((64, 1024), float(i)), b=torch.arange(10) + i, c=None) for i in range(5)]
    seen = 0
    for i, dev in enumerate(ip.DevicePrefetcher(host, DEV)):
        assert dev["a"].is_cuda and dev["c"] is None
        assert torch.equal(dev["a"].cpu(), host[i]["a"]) and torch.equal(dev["b"].cpu(), host[i]["b"])
        seen += 1
    assert seen == 5
    with pytest.raises(lib.VtError):
        ip.DevicePrefetcher(host, "cpu")


def test_bf16_feature_cache_gives_the_same_step():
    """Region features / location embeddings handed over in bf16 (a feature cache: half the H2D bytes) == the same values in fp32,
    bit for bit: the first thing the projection does with them is the bf16 cast."""
    from visitron_b200 import modeling, synthetic
    from visitron_b200.pretrained import BertConfig
    cfg = BertConfig(30522, num_hidden_layers=1, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    cfg.img_feature_dim, cfg.detector_classes, cfg.action_space = 2054, 1601, 36
    torch.manual_seed(0)
    model = modeling.PreTrainOscar(cfg).to(DEV).train()
    batch = {k: v.to(DEV) for k, v in synthetic.pretrain_batch(2, 5, text_len=64, regions=16).items()}
    batch["img_feats"] = batch["img_feats"].bfloat16().float()
    batch["img_location_embeddings"] = batch["img_location_embeddings"].bfloat16().float()
    res = []
    for cached in (False, True):
        b = dict(batch)
        if cached:
            b["img_feats"], b["img_location_embeddings"] = b["img_feats"].bfloat16(), b["img_location_embeddings"].bfloat16()
        model.zero_grad(set_to_none=True)
        out = model(**b)
        out[0].backward()
        res.append((out[0].detach().clone(), model.bert.img_embedding.weight.grad.detach().clone()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_new_kernels_stay_inside_their_buffers():
    """Guard bands around every output of the input-row kernels and of the key-tile plan (the pool has no compute-sanitizer): the C ABI is
    called with pointers into the middle of sentinel-filled buffers and the bands must come back untouched."""
    from visitron_b200 import ops
    L = lib.init()
    st = lib.stream_ptr()
    G = 4096   # guard elements on either side

    def banded(n, dtype, fill):
        buf = torch.full((n + 2 * G,), fill, dtype=dtype, device=DEV)
        return buf, buf[G:G + n]

    def intact(buf, n, fill):
        return bool((buf[:G] == fill).all()) and bool((buf[G + n:] == fill).all())

    # masking: n = 7 * 333 tokens (not a multiple of the block size)
    ids, tc = _batch(7, 333, 3, True)
    n = ids.numel()
    d_ids, d_tc, lut = ids.to(DEV), tc.to(DEV), ip.special_lut(V, SPECIAL, DEV)
    b_out, out = banded(n, torch.int64, -77)
    b_lab, lab = banded(n, torch.int64, -77)
    b_att, att = banded(n, torch.uint8, 99)
    lib.check(L.vt_mlm_mask(d_ids.data_ptr(), d_tc.data_ptr(), lut.data_ptr(), n, V, MASK, PAD, 0.15, 5, out.data_ptr(), lab.data_ptr(),
                            att.data_ptr(), st), "vt_mlm_mask")
    b_u, u = banded(3 * n, torch.float32, -5.0)
    b_r, r = banded(n, torch.int64, -77)
    lib.check(L.vt_mlm_draws(n, V, 5, u.data_ptr(), r.data_ptr(), st), "vt_mlm_draws")
    torch.cuda.synchronize()
    assert intact(b_out, n, -77) and intact(b_lab, n, -77) and intact(b_att, n, 99) and intact(b_u, 3 * n, -5.0) and intact(b_r, n, -77)
    assert bool((att <= 1).all()) and bool((u >= 0).all()) and bool((u < 1).all()) and int(r.min()) >= 0 and int(r.max()) < V

    # region packing: odd widths (scalar path) and the even widths of the model (vector path); ragged counts incl. 0 and > R
    for D, DL in ((7, 5), (2054, 128)):
        counts, R = [3, 0, 9, 40], 8
        feats, loc, off = _ragged(counts, D, DL, 1)
        B = len(counts)
        b_f, f = banded(B * R * D, torch.float32, -5.0)
        b_l, l_ = banded(B * R * DL, torch.float32, -5.0)
        b_a, a = banded(B * (2 + R), torch.int64, -77)
        lib.check(L.vt_pack_regions(feats.to(DEV).data_ptr(), D, loc.to(DEV).data_ptr(), DL, off.to(DEV).data_ptr(), B, R, f.data_ptr(), l_.data_ptr(),
                                    a.data_ptr() + 16, 2 + R, st), "vt_pack_regions")
        torch.cuda.synchronize()
        assert intact(b_f, B * R * D, -5.0) and intact(b_l, B * R * DL, -5.0) and intact(b_a, B * (2 + R), -77)
        av = a.view(B, 2 + R)
        assert bool((av[:, :2] == -77).all()), "the columns left of the attention tail belong to the caller"
        assert av[:, 2:].sum(1).tolist() == [min(c, R) for c in counts]

    # key-tile plan: B = 5 samples, S = 700 (11 tiles, the last one ragged)
    mask = torch.ones(5, 700, dtype=torch.long)
    mask[1, 100:] = 0
    bias = ops.mask_bias(mask.to(DEV))
    b_p, plan = banded(5, torch.int32, -77)
    lib.check(L.vt_attention_key_tiles(bias.data_ptr(), 5, 700, plan.data_ptr(), st), "vt_attention_key_tiles")
    torch.cuda.synchronize()
    assert intact(b_p, 5, -77) and plan.tolist() == [2047, 3, 2047, 2047, 2047]
