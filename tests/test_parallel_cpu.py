"""Data-parallel path on CPU: two gloo ranks, each on its own shard of the batch, must end the step with the
AVERAGE of the two ranks' gradients in every parameter's .grad (what DDP gives the reference loops,
vs/pretrain.py:96-102), bucket by bucket, and identical weights after an optimizer step.  Kernels are replaced by
the CPU test double (tests/abi_double.py); this checks the reducer / bucket / autograd plumbing only."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _Patch:
    def setattr(self, obj, name, val):
        setattr(obj, name, val)


def _build(seed=5):
    from oracle import visitron_oracle as vo
    from visitron_b200 import modeling
    from visitron_b200.pretrained import BertConfig
    oc = vo.OracleConfig(vocab_size=400, num_hidden_layers=2, max_position_embeddings=16, detector_classes=24)
    c = BertConfig(oc.vocab_size, num_hidden_layers=2, max_position_embeddings=16, hidden_dropout_prob=0.0,
                   attention_probs_dropout_prob=0.0)
    c.detector_classes = 24
    torch.manual_seed(seed)
    m = modeling.PreTrainOscar(c)
    m.train()
    return m


def _batch(rank):
    from visitron_b200 import synthetic
    return synthetic.pretrain_batch(2, 100 + rank, text_len=12, regions=8, vocab=400, detector_classes=24)


def _worker(rank, world, port, out_dir):
    sys.path[:0] = [os.path.dirname(HERE), HERE]
    import abi_double
    abi_double.install(_Patch())
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from visitron_b200 import parallel
        model = _build(seed=5 + rank)              # different init per rank: distribute() must broadcast rank 0's
        parallel.distribute(model)
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        out = model(**_batch(rank))
        loss = out[0]
        loss.backward()
        red = model._vt_reducer
        grads = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
        opt.step()
        scal = parallel.average_scalars(*out)
        torch.save(dict(grads=grads, calls=red.calls, bytes=red.bytes_reduced, loss=float(loss.detach()),
                        mean_loss=float(scal[0]), w=model.bert.pooler.dense.weight.detach().clone()),
                   os.path.join(out_dir, "rank%d.pt" % rank))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_gradient_average(tmp_path, monkeypatch):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    # single-process truth: rank 0's weights, each shard on its own, averaged
    import abi_double
    abi_double.install(monkeypatch)
    ref = {}
    losses = []
    for rank in range(2):
        m = _build(seed=5)
        out = m(**_batch(rank))
        out[0].backward()
        losses.append(float(out[0].detach()))
        for k, p in m.named_parameters():
            if p.grad is not None:
                ref[k] = ref.get(k, 0) + p.grad / 2
    assert set(r0["grads"]) == set(ref)
    for k in ref:
        torch.testing.assert_close(r0["grads"][k], ref[k], rtol=1e-4, atol=1e-7, msg=k)
        assert torch.equal(r0["grads"][k], r1["grads"][k]), k
    assert abs(r0["loss"] - losses[0]) < 1e-5 and abs(r1["loss"] - losses[1]) < 1e-5
    assert abs(r0["mean_loss"] - sum(losses) / 2) < 1e-5
    assert torch.equal(r0["w"], r1["w"])
    # heads + one bucket per layer + embeddings/image bucket
    assert r0["calls"] == 1 + 2 + 1
    n_params = sum(p.numel() for p in _build().parameters())
    assert r0["bytes"] >= 4 * n_params
