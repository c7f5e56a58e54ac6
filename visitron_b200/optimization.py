"""The step after backward: AdamW with the reference's semantics as ONE launch, fused
gradient-norm clipping, and the warm-up schedules.

Drop-in for ``transformers/pytorch_transformers/optimization.py`` as the loops use it
(vs/pretrain.py:109-139,192-193; vs/agent.py:514-518):

    optimizer = AdamW(optimizer_grouped_parameters, lr=args.learning_rate, eps=args.adam_epsilon)
    scheduler = WarmupLinearSchedule(optimizer, warmup_steps=..., t_total=...)
    ...
    loss.backward(); clip_grad_norm_(params, 40.0); optimizer.step(); scheduler.step()

Same constructor, same ``param_groups`` / ``state`` layout (``step``, ``exp_avg``, ``exp_avg_sq`` per parameter, so
optimizer checkpoints interchange), same arithmetic (optimization.py:165-187: decoupled decay applied AFTER the Adam
update with the un-corrected lr, eps outside the square root, optional bias correction).  The reference issues ~10 ATen
launches per parameter (200 parameters); here a device-resident segment table is built once and every parameter of
every group is updated by one kernel (visitron_b200/csrc/optim.cu).  CUDA fp32 parameters only: anything else raises,
there is no eager fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Iterable, List, Optional

import torch
from torch.optim import Optimizer
from torch.optim.lr_scheduler import LambdaLR

from . import lib

MAX_GROUPS = 8


class _Segment(C.Structure):  # vt_opt_segment
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p), ("p_bf16", C.c_void_p), ("n", C.c_int64),
                ("group", C.c_int32), ("first_chunk", C.c_int32)]


class _Table:
    """Segment table on the device for a list of (param, grad, m, v, group); rebuilt only when a pointer moves."""

    def __init__(self):
        self.key = None
        self.dev: Optional[torch.Tensor] = None
        self.n_segments = 0
        self.n_chunks = 0
        self.stores = {}

    def update(self, entries, device, with_shadow: bool = False) -> "_Table":
        base = tuple((p.data_ptr(), g.data_ptr(), 0 if m is None else m.data_ptr(), p.numel(), grp) for p, g, m, _, grp in entries)
        if with_shadow:
            from . import flat
            live = tuple((id(st), 0 if st.bf16 is None else st.bf16.data_ptr()) for st in list(flat._STORES))
        else:
            live = ()
        key = (base, live)     # the shadow pointers are a function of the parameter pointers and of the stores' bf16 buffers
        if key == self.key:
            return self
        shadows = [None] * len(entries)
        if with_shadow:
            shadows = [flat.shadow_of(e[0]) for e in entries]
        self.stores = {id(sh[0]): sh[0] for sh in shadows if sh is not None}
        chunk = int(lib.load().vt_opt_chunk())
        arr = (_Segment * len(entries))()
        first = 0
        for i, (p, g, m, v, grp) in enumerate(entries):
            arr[i].p, arr[i].g = p.data_ptr(), g.data_ptr()
            arr[i].m = 0 if m is None else m.data_ptr()
            arr[i].v = 0 if v is None else v.data_ptr()
            arr[i].p_bf16 = 0 if shadows[i] is None else shadows[i][1].data_ptr()
            arr[i].n, arr[i].group, arr[i].first_chunk = p.numel(), grp, first
            first += (p.numel() + chunk - 1) // chunk
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
        self.dev = raw.to(device)
        self.key, self.n_segments, self.n_chunks = key, len(entries), first
        return self


def _check_param(p: torch.Tensor, g: torch.Tensor):
    lib.require_device(p)
    if p.dtype != torch.float32 or g.dtype != torch.float32:
        raise lib.VtError("visitron_b200.optimization updates fp32 master weights with fp32 gradients (got %s / %s)" % (p.dtype, g.dtype))
    if g.is_sparse:
        raise RuntimeError("Adam does not support sparse gradients, please consider SparseAdam instead")
    if not p.is_contiguous() or not g.is_contiguous():
        raise lib.VtError("parameters and gradients must be contiguous")


class AdamW(Optimizer):
    """Adam with the weight-decay fix; constructor and semantics of optimization.py:107-189."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, correct_bias=True):
        for what, value, ok in (("learning rate", lr, lr >= 0.0), ("beta parameter", betas[0], 0.0 <= betas[0] < 1.0),
                                ("beta parameter", betas[1], 0.0 <= betas[1] < 1.0), ("epsilon value", eps, eps >= 0.0)):
            if not ok:   # same ValueError conditions as optimization.py:117-124
                raise ValueError("Invalid %s: %r" % (what, value))
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, correct_bias=correct_bias)
        super().__init__(params, defaults)
        self._table = _Table()
        self.grad_scale: Optional[torch.Tensor] = None   # device scalar applied to every gradient by the next step (see clip_grad_norm_)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # kernel launches are grouped by (betas, eps); the per-group lr / decay / step size travel as kernel scalars
        buckets = {}
        for group in self.param_groups:
            hp = (float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]))
            for p in group["params"]:
                if p.grad is None:
                    continue
                _check_param(p, p.grad)
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state["step"] += 1
                step_size = group["lr"]
                if group["correct_bias"]:   # "No bias correction for Bert"
                    bc1 = 1.0 - hp[0] ** state["step"]
                    bc2 = 1.0 - hp[1] ** state["step"]
                    step_size = step_size * math.sqrt(bc2) / bc1
                sub = buckets.setdefault(hp, {})
                sub.setdefault((float(step_size), float(group["lr"] * group["weight_decay"])), []).append(p)
        if not buckets:
            return loss
        if len(buckets) > 1:
            raise lib.VtError("all parameter groups must share betas and eps (the reference's loops do)")
        (b1, b2, eps), sub = next(iter(buckets.items()))
        if len(sub) > MAX_GROUPS:
            raise lib.VtError("at most %d distinct (step size, decay) groups per step" % MAX_GROUPS)
        entries, steps, decays = [], [], []
        for gi, ((step_size, lrd), plist) in enumerate(sub.items()):
            steps.append(step_size)
            decays.append(lrd)
            for p in plist:
                st = self.state[p]
                entries.append((p, p.grad, st["exp_avg"], st["exp_avg_sq"], gi))
        dev = entries[0][0].device
        with torch.cuda.device(dev):
            t = self._table.update(entries, dev, with_shadow=True)
            fs = (C.c_float * len(steps))(*steps)
            fd = (C.c_float * len(decays))(*decays)
            lib.check(lib.init(dev.index).vt_adamw_step(t.dev.data_ptr(), t.n_segments, t.n_chunks, len(steps), fs, fd, b1, b2, eps,
                                                        lib.ptr(self.grad_scale), lib.stream_ptr()), "vt_adamw_step")
        self.grad_scale = None
        # the kernel wrote the parameters behind autograd's back: bump their version counters, so that saved-tensor checks
        # and the bf16 operand cache (flat.FlatStore.refresh_bf16) see the update
        torch.autograd.graph.increment_version([e[0] for e in entries])
        for st in t.stores.values():   # the launch wrote the bf16 operand of every weight it updated: the next forward skips its cast
            st.mark_fresh()
        return loss


class _ClipState:
    table = _Table()
    partials: Optional[torch.Tensor] = None
    out: Optional[torch.Tensor] = None


def clip_grad_norm_(parameters: Iterable[torch.Tensor], max_norm: float, norm_type: float = 2.0, optimizer: Optional[AdamW] = None) -> torch.Tensor:
    """``torch.nn.utils.clip_grad_norm_`` (vs/agent.py:514-518) in 3 launches for any number of parameters: per-chunk sums
    of squares, a deterministic final reduction producing (norm, coefficient), and the in-place scale (which returns
    immediately when the norm is within the bound).  With ``optimizer=`` the scale is folded into that optimizer's next
    ``step()`` instead of touching the gradients.  Returns the total norm as a 0-d device tensor (no host sync)."""
    if norm_type != 2.0:
        raise NotImplementedError("only the L2 norm is used by the reference's loops")
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    ps: List[torch.Tensor] = [p for p in parameters if p.grad is not None]
    if not ps:
        return torch.zeros(())
    for p in ps:
        _check_param(p, p.grad)
    dev = ps[0].device
    with torch.cuda.device(dev):
        t = _ClipState.table.update([(p, p.grad, None, None, 0) for p in ps], dev)
        if _ClipState.partials is None or _ClipState.partials.numel() < t.n_chunks or _ClipState.partials.device != dev:
            _ClipState.partials = torch.empty(t.n_chunks, dtype=torch.float32, device=dev)
        out = torch.empty(2, dtype=torch.float32, device=dev)
        lib.check(lib.init(dev.index).vt_clip_grad_norm(t.dev.data_ptr(), t.n_segments, t.n_chunks, float(max_norm), _ClipState.partials.data_ptr(),
                                                        out.data_ptr(), 0 if optimizer is not None else 1, lib.stream_ptr()), "vt_clip_grad_norm")
    if optimizer is not None:
        optimizer.grad_scale = out[1:2]
    else:
        torch.autograd.graph.increment_version([p.grad for p in ps])
    return out[0]


# ---- schedules (optimization.py:26-63): host-side scalars --------------------------------------------------------------
class _RampThen(LambdaLR):
    """lr multiplier = step / warmup_steps while warming up, then ``self.after(step)``."""

    def __init__(self, optimizer, warmup_steps, last_epoch=-1):
        self.warmup_steps = warmup_steps
        super().__init__(optimizer, self.lr_lambda, last_epoch=last_epoch)

    def after(self, step) -> float:
        return 1.0

    def lr_lambda(self, step):
        ramping = step < self.warmup_steps
        return step / max(1.0, float(self.warmup_steps)) if ramping else self.after(step)


class ConstantLRSchedule(_RampThen):
    def __init__(self, optimizer, last_epoch=-1):
        super().__init__(optimizer, 0, last_epoch)


class WarmupConstantSchedule(_RampThen):
    """``--scheduler constant`` (vs/pretrain.py:132-133)."""


class WarmupLinearSchedule(_RampThen):
    """``--scheduler linear`` (vs/pretrain.py:134-137): after the ramp, a straight line down to 0 at ``t_total``."""

    def __init__(self, optimizer, warmup_steps, t_total, last_epoch=-1):
        self.t_total = t_total
        super().__init__(optimizer, warmup_steps, last_epoch)

    def after(self, step) -> float:
        left = float(self.t_total - step)
        return max(0.0, left / max(1.0, float(self.t_total - self.warmup_steps)))
