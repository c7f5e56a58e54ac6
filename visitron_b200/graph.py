"""One training step (forward + backward) as ONE CUDA graph.

The reference's loop (vs/pretrain.py:148-201) issues ~1500 eager launches per step and syncs at its seven ``.item()``
reads; round 1 of this library issued ~400 launches through ctypes and synced once per step (``torch.nonzero`` on the
labels).  Both leave the GPU idle whenever the host thread is descheduled, and at the scripts' batch of 2 per GPU the
step is bound by the host's launch rate on any GPU.  Here the step has a FIXED launch sequence -- the labelled rows of the
pre-training heads are compacted and counted on the device (``ops.compact_labels`` / ``ops.row_limit``), the dropout seed
lives in device memory and is advanced by the graph's first node (``ops.seed_advance`` / ``ops.set_seed_source``) -- so the
whole step is captured once and replayed with one ``cudaGraphLaunch``; the host can queue many steps ahead.

    step = GraphedStep(model, example_batch)        # device tensors of the shapes every batch will have; captured once
    for batch in loader:
        loss, *scalars = step(batch)                # copies the batch into the static inputs, replays the graph
        optimizer.step(); scheduler.step()          # p.grad are the static gradient views the replay has just filled

The gradients land in the same flat buffers as in eager mode (``flat.FlatStore``), the data-parallel exchange of
``parallel.distribute`` is captured with the step (NCCL collectives are capturable), and eager calls of the model keep
working beside the graph (e.g. validation).  Shapes are baked in: a batch of another shape needs another GraphedStep.
"""
from __future__ import annotations

import gc
from typing import Callable, Dict, Iterable, Optional, Sequence

import torch

from . import lib, modeling, ops


def _default_step(model, batch):
    return model(**batch)


class GraphedStep:
    def __init__(self, model: torch.nn.Module, example_batch: Dict[str, torch.Tensor],
                 step_fn: Optional[Callable] = None, extra_params: Iterable[torch.nn.Parameter] = (), warmup: int = 2,
                 seed: Optional[int] = None, post_backward: Optional[Callable[[], None]] = None):
        """``step_fn(model, batch)`` returns a tuple of 0-d tensors whose first element is the loss (default: ``model(**batch)``,
        the call of vs/pretrain.py:167).  ``extra_params``: parameters outside ``model`` that the step trains (an agent's heads).
        ``post_backward``: captured behind the backward (e.g. the all-reduce of extra parameters)."""
        self.model = model
        self.fn = step_fn or _default_step
        self.params = list(model.parameters()) + list(extra_params)
        dev = next(model.parameters()).device
        lib.require_device(next(model.parameters()))
        self.device = dev
        for k, v in example_batch.items():
            if not (isinstance(v, torch.Tensor) and v.is_cuda):
                raise lib.VtError("GraphedStep needs device tensors in the example batch (got %s for %r)" % (type(v).__name__, k))
        self.static_in = {k: v.clone() for k, v in example_batch.items()}
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())   # CPU generator: reproducible under torch.manual_seed
            if torch.distributed.is_available() and torch.distributed.is_initialized():
                seed ^= (torch.distributed.get_rank() + 1) * 0x9E3779B97F4A7C15 & (2 ** 63 - 1)   # ranks must not share masks
        self.seed_state = torch.tensor([seed, 0], dtype=torch.int64, device=dev)   # (counter, seed of the step being run)
        # eager warm-up on a side stream: lazy initialisation (function attributes, NCCL communicators, allocator growth)
        # must not happen inside the capture
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._clear_grads()
                out = self.fn(model, self.static_in)
                out[0].backward()
                if post_backward is not None:
                    post_backward()
                del out      # a live loss keeps this iteration's AccumulateGrad nodes (and their stream) alive into the capture
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._clear_grads()
        # no autograd graph of an earlier iteration may survive into the capture: its AccumulateGrad nodes carry the stream they
        # were created on, and the engine's end-of-backward sync with that stream would invalidate the capture
        gc.collect()
        self.graph = torch.cuda.CUDAGraph()
        n0 = lib.launch_count()
        modeling._CAPTURING = True
        ops.set_seed_source(self.seed_state)
        try:
            with torch.cuda.graph(self.graph):
                ops.seed_advance(self.seed_state)
                out = self.fn(model, self.static_in)
                out[0].backward()
                if post_backward is not None:
                    post_backward()
                self.static_out = torch.stack([o.detach().float().reshape(()) for o in out])
        finally:
            modeling._CAPTURING = False
            ops.set_seed_source(None)
        self.launches_per_step = lib.launch_count() - n0     # kernels of libvisitron_b200.so inside one replay
        self.n_out = len(out)
        self._grads = [(p, p.grad) for p in self.params if p.grad is not None]

    def _clear_grads(self):
        for p in self.params:
            p.grad = None

    def load(self, batch: Dict[str, torch.Tensor]) -> None:
        """Copies a batch (device tensors, or pinned host tensors) into the graph's static inputs on the current stream."""
        for k, dst in self.static_in.items():
            src = batch[k]
            if src.shape != dst.shape or src.dtype != dst.dtype:
                raise lib.VtError("GraphedStep was captured for %r of shape %s / %s, got %s / %s"
                                  % (k, tuple(dst.shape), dst.dtype, tuple(src.shape), src.dtype))
            if src.data_ptr() != dst.data_ptr():
                dst.copy_(src, non_blocking=True)

    def replay(self) -> torch.Tensor:
        """One step on the static inputs; returns the step's scalars as ONE device tensor [n_out] (overwritten by the next replay)."""
        self.graph.replay()
        for p, g in self._grads:      # zero_grad(set_to_none=True) between steps must not detach the static gradient views
            if p.grad is not g:
                p.grad = g
        return self.static_out

    def __call__(self, batch: Optional[Dict[str, torch.Tensor]] = None) -> Sequence[torch.Tensor]:
        if batch is not None:
            self.load(batch)
        return self.replay().unbind(0)
