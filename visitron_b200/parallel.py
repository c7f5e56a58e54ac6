"""Data-parallel gradient exchange: bucketed all-reduce (average) overlapped with backward.

Replaces the reference's ``DistributedDataParallel(model, find_unused_parameters=True)`` wrap
(vs/pretrain.py:96-102, vs/train.py:98-110).  The batch is sharded over one process per GPU; the
only exchange step of an iteration is the gradient average.  Gradients are produced directly inside
contiguous slices of the flat gradient buffer (visitron_b200/flat.py), one slice per encoder layer
(~28 MB fp32, DDP's own bucket size is 25 MB), in reverse execution order: heads, pooler + layer 11,
... layer 0, image projection + embeddings.  As soon as the backward kernels of a slice have been
enqueued, its all-reduce is enqueued on NCCL's stream (over NVLink 5 / NVSwitch), so it runs
under the next layer's backward GEMMs; the autograd node returns only after the current stream
has been made to wait for all of them (a stream-level wait, the host does not block).

Usage (reference loop, unchanged otherwise):

    model = PreTrainOscar.from_pretrained(...).to(device)
    visitron_b200.parallel.distribute(model)          # instead of DDP(model, ...)

Like DDP the reducer AVERAGES (the loops divide the loss by world size themselves as
well; that quirk is theirs and is preserved by not compensating for it here).
Parameters a call does not touch (fine-tuning leaves img_embedding / location_embeds / pooler
unused) are skipped consistently on every rank: which buckets exist depends only on the call
signature, never on data.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist

# CTAs (= SMs) NCCL may use for the overlapped gradient exchange; the persistent GEMM leaves as many SMs free while
# an exchange is in flight, so that neither side waits for the other's CTAs to retire.  NVSwitch reduces in the switch
# (NVLS), so a 28 MB bucket needs few CTAs to run far faster than the backward of the next layer.  Measured at b = 32 per GPU
# (ms per step, cap 2 / 4 / 8 / 16): 2 GPUs 25.0 / 24.7 / 26.5 / 25.9; 8 GPUs - / 25.3 / 26.7 / - (1 GPU: 23.7-24.1).
# Round 2 (17 ms step, `tools/r2_call30.sh` / `r2_call31.sh`): 2 GPUs 17.87 / 18.38 ms with 4 / 8 CTAs, 8 GPUs 19.33 / 18.69 / 18.99 ms with 4 / 8 / 16
# -- an 8-rank all-reduce moves 1.75x the bytes of a 2-rank one per GPU and falls behind the backward with 4 CTAs.  Hence 4 CTAs
# up to 3 ranks and 8 from 4 ranks on (VT_NCCL_MAX_CTAS overrides).  Skipping the step's last (101 MB) all-reduce altogether gains
# 0.0 ms at 2 GPUs and 0.34 ms at 8: what is left above the 1-GPU step is the overlapped exchange's share of SMs, HBM and power.
NCCL_MAX_CTAS = int(os.environ.get("VT_NCCL_MAX_CTAS", "0"))


def _max_ctas(world_size: int) -> int:
    return NCCL_MAX_CTAS if NCCL_MAX_CTAS > 0 else (4 if world_size < 4 else 8)


_DBG_SKIP_TAIL = os.environ.get("VT_DBG_SKIP_TAIL", "0") == "1"
_DBG_NO_RESERVE = os.environ.get("VT_DBG_NO_RESERVE", "0") == "1"


def _exchange_group(max_ctas: int):
    """A NCCL communicator of its own for the gradient exchange, capped at `max_ctas` CTAs."""
    try:
        opts = dist.ProcessGroupNCCL.Options()
        opts.config.max_ctas = max_ctas
        opts.config.min_ctas = 1
        return dist.new_group(backend="nccl", pg_options=opts)
    except Exception:  # older torch without ncclConfig plumbing: the default communicator
        return None


class GradReducer:
    def __init__(self, process_group=None, overlap: bool = True):
        if not dist.is_available() or not dist.is_initialized():
            raise RuntimeError("GradReducer needs an initialised torch.distributed process group")
        self.backend = dist.get_backend(process_group)
        self.world_size = dist.get_world_size(process_group)
        self.reserve = 0
        caller_group = process_group
        if self.backend == "nccl" and process_group is None and overlap and self.world_size > 1:
            process_group = _exchange_group(_max_ctas(self.world_size))
            self.reserve = _max_ctas(self.world_size) if process_group is not None else 0
        # the LAST bucket of a step (embeddings, 101 MB) has no backward kernels left to hide behind: it goes through the
        # caller's (uncapped) communicator at full speed instead of the 8-CTA one
        self.tail_group = caller_group
        self.tail_full_speed = os.environ.get("VT_TAIL_FULL_SPEED", "1") != "0"
        self.group = process_group
        self._reserved = False
        self.overlap = overlap
        self._works: List = []
        self._pending: List[torch.Tensor] = []
        self.bytes_reduced = 0
        self.calls = 0

    def reduce(self, bucket: torch.Tensor, tail: bool = False) -> None:
        """Enqueue the average of ``bucket`` (a contiguous slice of the flat gradient buffer); ``tail``: the step's last."""
        if self.world_size == 1 or bucket.numel() == 0:
            return
        self.calls += 1
        self.bytes_reduced += bucket.numel() * bucket.element_size()
        if tail and _DBG_SKIP_TAIL:      # timing ablation only (wrong gradients): what the step's last all-reduce costs
            return
        if tail and self.tail_full_speed and self.reserve and self.backend == "nccl":
            self._works.append(dist.all_reduce(bucket, op=dist.ReduceOp.AVG, group=self.tail_group, async_op=True))
            return
        if self.reserve and not self._reserved and not _DBG_NO_RESERVE:
            from . import lib
            lib.load().vt_reserve_sms(self.reserve)
            self._reserved = True
        if self.backend == "nccl":
            w = dist.all_reduce(bucket, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
        else:  # gloo (CPU tests): no AVG
            w = dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._pending.append(bucket)
        self._works.append(w)
        if not self.overlap:
            self.finish()

    def finish(self) -> None:
        """Make the current stream wait for every enqueued all-reduce (host does not block on NCCL)."""
        for w in self._works:
            w.wait()
        self._works = []
        if self._reserved:
            from . import lib
            lib.load().vt_reserve_sms(0)
            self._reserved = False
        for b in self._pending:
            b.div_(self.world_size)
        self._pending = []


def distribute(model, process_group=None, overlap: bool = True):
    """Attach a GradReducer to a visitron_b200 model (and its ``.bert`` encoder): the drop-in
    for ``DistributedDataParallel(model, ...)``; returns the model itself (no ``.module`` wrapper,
    ``hasattr(model, 'module')`` checks in the reference's save code, vs/pretrain.py:258-260, stay false)."""
    reducer = GradReducer(process_group, overlap=overlap)
    n = 0
    for m in model.modules():
        if hasattr(m, "_vt_set_reducer"):
            m._vt_set_reducer(reducer)
            n += 1
    if n == 0:
        raise TypeError("distribute() expects a visitron_b200 model (BertImgModelwithLocationEmbeds / PreTrainOscar / OscarEncoder)")
    broadcast_parameters(model, process_group)
    return model


def broadcast_parameters(model, process_group=None, src: int = 0) -> None:
    """Rank ``src``'s weights everywhere (what DDP's constructor does)."""
    with torch.no_grad():
        for p in model.parameters():
            dist.broadcast(p.data, src=src, group=process_group)
        for b in model.buffers():
            dist.broadcast(b.data, src=src, group=process_group)


def average_scalars(*values: torch.Tensor, process_group=None) -> List[torch.Tensor]:
    """The loop's 7 scalar all-reduces (vs/pretrain.py:169-189) as ONE 7-float all-reduce."""
    if not dist.is_initialized() or dist.get_world_size(process_group) == 1:
        return list(values)
    buf = torch.stack([v.detach().float().reshape(()) for v in values])
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=process_group)
    buf /= dist.get_world_size(process_group)
    return list(buf.unbind(0))
