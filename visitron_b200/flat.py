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

import weakref
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import ops

ALIGN = 8  # elements: 32 B in fp32, 16 B in bf16 (TMA / 128-bit vector alignment)

_STORES: "weakref.WeakSet[FlatStore]" = weakref.WeakSet()   # every live store: the fused AdamW looks up a parameter's bf16 shadow here


def shadow_of(p: torch.Tensor):
    """(store, bf16 view with p's shape) when `p` is a view into a FlatStore whose bf16 operand buffer exists, else None:
    visitron_b200.optimization.AdamW writes the updated weight there in its own launch (no separate cast pass)."""
    a = p.data_ptr()
    for st in list(_STORES):
        if st.flat is None or st.bf16 is None:
            continue
        base = st.flat.data_ptr()
        if base <= a < base + 4 * st.total:
            off = (a - base) // 4
            return st, st.bf16[off: off + p.numel()]
    return None


def invalidate(model=None) -> None:
    """Forget every 'the bf16 operands are fresh' note (call after editing weights through `p.data` right behind
    visitron_b200.optimization.AdamW.step(); every other way of changing a weight is noticed without help)."""
    for st in list(_STORES):
        st._fresh = None


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
        self._fresh = None     # version tuple behind which the fused AdamW vouched for the bf16 copy (single use)
        self._cur = 0
        _STORES.add(self)
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
        self._fresh = None
        self._wc, self._wbc, self._gc = {}, {}, [{}, {}]
        return self

    def refresh_bf16(self) -> None:
        """The bf16 tensor-core operands of this forward: ONE cast launch over the flat fp32 masters, at the start of EVERY
        forward (0.68 GB of traffic, ~0.12 ms for the 113 M parameters, inside the timed step).

        Round 1 skipped the cast while the parameters' autograd version counters stood still.  In-place updates through
        ``p.data`` -- the reference's own AdamW (transformers/pytorch_transformers/optimization.py:176,187), which the loops
        import unmodified, or ``dist.broadcast(p.data)`` -- do not move those counters, so the GEMMs kept multiplying by the
        step-0 weights.  Nothing the host can observe proves that the masters are unchanged, hence no cache.  The one exception
        is an update made by THIS library: visitron_b200.optimization.AdamW writes the bf16 copy of every weight it updates in
        the same launch and leaves a single-use note (``mark_fresh``) that is honoured only if every version counter still has
        the value the optimizer left behind."""
        if self.bf16 is None:
            self.bf16 = torch.empty(self.total, dtype=torch.bfloat16, device=self.flat.device)
            self._wbc = {}
            self._fresh = None
        fresh, self._fresh = self._fresh, None
        if fresh is not None and fresh == tuple(p._version for p in self.params):
            return
        ops.to_bf16(self.flat, out=self.bf16)

    def mark_fresh(self) -> None:
        """Called by the fused AdamW after it has written the bf16 copies of the weights it updated (the others are unchanged
        since the last forward's cast)."""
        if self.bf16 is not None:
            self._fresh = tuple(p._version for p in self.params)

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
        live = [p.grad.data_ptr() for p in self.params if p.grad is not None]

        def aliased(buf) -> bool:
            lo = buf.data_ptr()
            hi = lo + 4 * self.total
            return any(lo <= a < hi for a in live)

        k = None
        for i, buf in enumerate(self._grad):
            if buf is None or not aliased(buf):
                k = i
                break
        if k is None:
            # both buffers still back a live .grad (a parameter that got its first gradient in a later step, with in-place
            # zeroing, ends up in the other buffer than its siblings): zeroing either would wipe an accumulated gradient
            self._grad.append(None)
            self._gc.append({})
            k = len(self._grad) - 1
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


    def assign_grads(self, used: Optional[set] = None) -> None:
        """Hands the gradient views to the parameters WITHOUT autograd's AccumulateGrad nodes (used while a step is captured in a
        CUDA graph: a node that outlived an earlier iteration carries that iteration's stream, and the engine's stream sync
        would invalidate the capture).  Same semantics: adopt the view, or add it to a gradient that is already there."""
        for view, p in zip(self.grad_views(used), self.params):
            if view is None:
                continue
            if p.grad is None:
                p.grad = view
            else:
                p.grad.add_(view)


class _LayoutChanged(Exception):
    pass
