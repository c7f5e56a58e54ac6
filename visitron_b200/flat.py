"""Flat parameter / gradient / bf16-operand storage.

The reference keeps 200 separately allocated fp32 ``nn.Parameter``s and lets DDP copy their
gradients into 25 MB buckets (vs/pretrain.py:96-102).  Here every parameter of a model is a VIEW
into one fp32 buffer laid out in execution order, with

  * a same-layout bf16 buffer refreshed by ONE cast launch per optimizer step (the tensor-core
    operands; query/key/value weights are adjacent, so the fused QKV projection reads one
    [2304, 768] block and its weight-gradient GEMM writes one [2304, 768] block whose thirds are
    the three parameters' gradients), and
  * a same-layout fp32 gradient buffer the backward kernels accumulate into directly; its
    contiguous slices are the all-reduce buckets (no gradient copies, HBM traffic = one write).

``state_dict`` keys, shapes and values are untouched: parameters stay ordinary fp32
``nn.Parameter``s registered on the same sub-modules as in the reference.
``model.to()`` / ``resize_specific_embeddings`` / ``deepcopy`` re-point parameters at fresh
storage; ``ensure()`` detects that and re-flattens.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import ops

ALIGN = 8  # elements: 32 B in fp32, 16 B in bf16 (TMA / 128-bit vector alignment)


def _round_up(n: int, m: int = ALIGN) -> int:
    return (n + m - 1) // m * m


class FlatStore:
    def __init__(self, named_params: Sequence[Tuple[str, nn.Parameter]], bucket_starts: Sequence[str] = ()):
        """``named_params``: layout order (duplicates of one Parameter object are laid out once).
        ``bucket_starts``: names at which a new gradient bucket begins."""
        self.names: List[str] = []
        self.params: List[nn.Parameter] = []
        self.offset: Dict[str, int] = {}
        self.numel: Dict[str, int] = {}
        self.shape: Dict[str, Tuple[int, ...]] = {}
        seen = {}
        off = 0
        for name, p in named_params:
            if id(p) in seen:
                raise ValueError("parameter %s laid out twice (also as %s)" % (name, seen[id(p)]))
            seen[id(p)] = name
            self.names.append(name)
            self.params.append(p)
            self.offset[name] = off
            self.numel[name] = p.numel()
            self.shape[name] = tuple(p.shape)
            off += _round_up(p.numel())
        self.total = off
        starts = sorted({self.offset[n] for n in bucket_starts} | {0})
        self.buckets: List[Tuple[int, int]] = [(s, e) for s, e in zip(starts, starts[1:] + [self.total]) if e > s]
        self.flat: Optional[torch.Tensor] = None
        self.bf16: Optional[torch.Tensor] = None
        self._grad: List[Optional[torch.Tensor]] = [None, None]
        self._versions = None
        self._cur = 0
        # view caches (a step asks for ~400 parameter / operand / gradient views; slicing + view cost ~4 us each on the host)
        self._wc: Dict = {}
        self._wbc: Dict = {}
        self._gc: List[Dict] = [{}, {}]

    # ------------------------------------------------------------------------------------
    def _aliased(self) -> bool:
        if self.flat is None:
            return False
        base = self.flat.data_ptr()
        for name, p in zip(self.names, self.params):
            if p.data_ptr() != base + 4 * self.offset[name] or tuple(p.shape) != self.shape[name]:
                return False
        return True

    def ensure(self) -> "FlatStore":
        """(Re-)points every parameter at the flat buffer; cheap when nothing moved."""
        if self._aliased():
            return self
        p0 = self.params[0]
        dev = p0.device
        for name, p in zip(self.names, self.params):
            if p.dtype != torch.float32:
                raise NotImplementedError("visitron_b200 keeps fp32 master weights; parameter %s is %s" % (name, p.dtype))
            if p.device != dev:
                raise ValueError("parameters live on different devices")
            if tuple(p.shape) != self.shape[name]:  # resized embedding table: a new layout is needed
                raise _LayoutChanged(name)
        flat = torch.zeros(self.total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for name, p in zip(self.names, self.params):
                view = flat[self.offset[name]: self.offset[name] + self.numel[name]].view(self.shape[name])
                view.copy_(p.data)
                p.data = view
        self.flat = flat
        self.bf16 = None
        self._grad = [None, None]
        self._versions = None
        self._wc, self._wbc, self._gc = {}, {}, [{}, {}]
        return self

    def refresh_bf16(self) -> None:
        """One cast launch per parameter update (keyed on the parameters' autograd version counters)."""
        vers = sum(p._version for p in self.params)
        if self.bf16 is not None and vers == self._versions:
            return
        if self.bf16 is None:
            self.bf16 = torch.empty(self.total, dtype=torch.bfloat16, device=self.flat.device)
            self._wbc = {}
        ops.to_bf16(self.flat, out=self.bf16)
        self._versions = vers

    # ------------------------------------------------------------------------------------
    def w(self, name: str) -> torch.Tensor:
        t = self._wc.get(name)
        if t is None:
            o = self.offset[name]
            t = self._wc[name] = self.flat[o: o + self.numel[name]].view(self.shape[name])
        return t

    def wb(self, name: str, rows: Optional[int] = None, cols: Optional[int] = None) -> torch.Tensor:
        """bf16 operand view; (rows, cols) overrides the shape to span adjacent parameters."""
        key = (name, rows, cols)
        t = self._wbc.get(key)
        if t is None:
            o = self.offset[name]
            if rows is None:
                t = self.bf16[o: o + self.numel[name]].view(self.shape[name])
            else:
                t = self.bf16[o: o + rows * cols].view(rows, cols)
            self._wbc[key] = t
        return t

    def wspan(self, name: str, n: int) -> torch.Tensor:
        o = self.offset[name]
        return self.flat[o: o + n]

    def g(self, name: str, rows: Optional[int] = None, cols: Optional[int] = None) -> torch.Tensor:
        cache = self._gc[self._cur]
        key = (name, rows, cols)
        t = cache.get(key)
        if t is None:
            buf = self._grad[self._cur]
            o = self.offset[name]
            if rows is None:
                t = buf[o: o + self.numel[name]].view(self.shape[name])
            elif cols is None:
                t = buf[o: o + rows]
            else:
                t = buf[o: o + rows * cols].view(rows, cols)
            cache[key] = t
        return t

    def grad_bucket(self, i: int) -> torch.Tensor:
        s, e = self.buckets[i]
        return self._grad[self._cur][s:e]

    def bucket_of(self, name: str) -> int:
        o = self.offset[name]
        for i, (s, e) in enumerate(self.buckets):
            if s <= o < e:
                return i
        raise KeyError(name)

    def begin_backward(self) -> None:
        """Selects (and zeroes) a gradient buffer no live ``.grad`` aliases: with
        ``zero_grad(set_to_none=True)`` that is always buffer 0; with in-place zeroing or gradient
        accumulation the views handed out last step are still the parameters' ``.grad`` and
        autograd will ``+=`` the new views into them, so the other buffer is used."""
        k = 0
        if self._grad[0] is not None:
            lo, hi = self._grad[0].data_ptr(), self._grad[0].data_ptr() + 4 * self.total
            for p in self.params:
                if p.grad is not None and lo <= p.grad.data_ptr() < hi:
                    k = 1
                    break
        if self._grad[k] is None:
            self._grad[k] = torch.empty(self.total, dtype=torch.float32, device=self.flat.device)
            self._gc[k] = {}
        self._grad[k].zero_()
        self._cur = k

    def grad_views(self, used: Optional[set] = None) -> List[Optional[torch.Tensor]]:
        """Fresh views (no other Python reference, so autograd adopts them without a copy) in
        ``self.params`` order; ``None`` for parameters outside ``used`` or not requiring grad."""
        out = []
        for name, p in zip(self.names, self.params):
            if not p.requires_grad or (used is not None and name not in used):
                out.append(None)
            else:   # NOT through the cache of g(): a cached view is a second reference and AccumulateGrad would copy
                o = self.offset[name]
                out.append(self._grad[self._cur][o: o + self.numel[name]].view(self.shape[name]))
        return out


class _LayoutChanged(Exception):
    pass
