"""ctypes binding of libvisitron_b200.so (the C ABI in include/visitron_b200.h).

PyTorch is used for device memory and streams only; every op here hands raw device pointers and
the current CUDA stream to the hand-written sm_100a kernels.  There is NO fallback: if the shared
library is missing or the device is not sm_100 every call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvisitron_b200.so")

ACT_NONE, ACT_GELU, ACT_TANH, ACT_GELU_BWD, ACT_TANH_BWD, ACT_GELU_SAVE_GRAD, ACT_MUL_AUX = 0, 1, 2, 3, 4, 5, 6


class VtError(RuntimeError):
    pass


class GemmEpilogue(C.Structure):
    _fields_ = [
        ("out_fp32", C.c_int), ("accumulate", C.c_int), ("split_k", C.c_int), ("act", C.c_int),
        ("alpha", C.c_float), ("alpha_dev", C.c_void_p), ("bias", C.c_void_p),
        ("residual", C.c_void_p), ("ldr", C.c_int64),
        ("aux", C.c_void_p), ("ldaux", C.c_int64),
        ("aux_out", C.c_void_p), ("ldaux_out", C.c_int64),
        ("dropout_p", C.c_float), ("seed", C.c_uint64), ("dropout_stream", C.c_uint32),
        ("remap_rows_in", C.c_int64), ("remap_rows_out", C.c_int64), ("remap_row_off", C.c_int64),
        ("remap_index", C.c_void_p),
        ("colsum", C.c_void_p),
    ]


_lib = None
_inited = set()
# bench.py's roofline pass: when a list, every GEMM launch is bracketed by CUDA events on the launching stream
# and (start, end, 2*M*N*K) is appended
GEMM_EVENTS = None
# the row-limit block (counts tensor, which) that ops.row_limit has set on this thread, for the roofline pass only: the rows a GEMM
# really walks are known on the device alone
ACTIVE_LIMIT = threading.local()


def load() -> C.CDLL:
    """Loads the shared library (no device needed); raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VtError(
            "libvisitron_b200.so not found at %s -- build it with `python -m visitron_b200.build`; "
            "there is no CPU / eager fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.vt_last_error.restype = C.c_char_p
    lib.vt_launch_count.restype = C.c_int64
    from . import _signatures
    _signatures.declare(lib)
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        raise VtError("%s failed (%d): %s" % (what, rc, load().vt_last_error().decode()))


_cur_dev = torch._C._cuda_getDevice if hasattr(torch._C, "_cuda_getDevice") else None
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def init(device: Optional[int] = None) -> C.CDLL:
    """The library handle, initialised for `device` (default: the current one).  On the hot path this is two dictionary /
    C lookups: a step makes ~400 calls, and `torch.cuda.is_available()` alone cost 2 ms per step there."""
    if _lib is not None and _inited:
        dev = (_cur_dev() if _cur_dev is not None else torch.cuda.current_device()) if device is None else int(device)
        if dev in _inited:
            return _lib
    lib = load()
    if not torch.cuda.is_available():
        raise VtError("visitron_b200 needs a CUDA sm_100 device; there is no CPU fallback")
    dev = torch.cuda.current_device() if device is None else int(device)
    if dev not in _inited:
        check(lib.vt_init(dev), "vt_init")
        _inited.add(dev)
    return lib


def require_device(t: torch.Tensor) -> None:
    """Product entry points call this first: CPU tensors are refused, never emulated."""
    if not t.is_cuda:
        raise VtError("visitron_b200 runs on a CUDA sm_100 device only; got a %s tensor (there is no CPU fallback)" % t.device)
    init(t.device.index)


def stream_ptr() -> int:
    """Raw handle of the current stream of the current device (the private fast path costs ~0.3 us, `torch.cuda.current_stream()`
    builds a Stream object every call: 14 us x 280 launches per step)."""
    if _raw_stream is not None:
        return _raw_stream(_cur_dev() if _cur_dev is not None else torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def launch_count() -> int:
    return int(load().vt_launch_count())


def reset_launch_count():
    load().vt_reset_launch_count()


def _row_major_2d(t: torch.Tensor, name: str):
    st = t.stride()
    if len(st) != 2 or st[1] != 1:
        raise VtError("%s must be a 2-D tensor with unit inner stride, got shape %s strides %s"
                      % (name, tuple(t.shape), t.stride()))
    return st[0] if t.size(0) > 1 else max(st[0], t.size(1))


_TLS = threading.local()   # one epilogue struct per thread, re-filled per call (a new ctypes Structure per GEMM cost ~8 us x 160 GEMMs per step)


def gemm(A: torch.Tensor, B: torch.Tensor, *, a_major: int = 0, b_major: int = 0,
         out: Optional[torch.Tensor] = None, out_dtype=torch.bfloat16, bias=None, act: int = ACT_NONE,
         alpha: float = 1.0, alpha_dev=None, residual=None, aux=None, aux_out=None,
         dropout_p: float = 0.0, seed: int = 0, dropout_stream: int = 0, accumulate: bool = False,
         split_k: int = 0, remap=None, remap_index=None, colsum=None, M=None, N=None, K=None) -> torch.Tensor:
    """D = epilogue(alpha * A x B^T) on the tcgen05 GEMM (see vt_gemm_bf16 in the header).

    a_major/b_major = 0: A is [M,K] / B is [N,K]; = 1: A is [K,M] / B is [K,N] (row-major).
    colsum (fp32 [M], a_major = 1 only): += alpha * column sums of A -- the bias gradient next to the weight gradient.
    """
    lib = init()
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    lda = _row_major_2d(A, "A")
    ldb = _row_major_2d(B, "B")
    m, ka = (A.size(1), A.size(0)) if a_major else (A.size(0), A.size(1))
    n, kb = (B.size(1), B.size(0)) if b_major else (B.size(0), B.size(1))
    M = m if M is None else M
    N = n if N is None else N
    K = ka if K is None else K
    if K != kb and (K > ka or K > kb):
        raise VtError("contraction mismatch: A has %d, B has %d" % (ka, kb))
    if out is None:
        if remap is not None or remap_index is not None:
            raise VtError("remap needs an explicit output tensor")
        out = torch.empty((M, N), device=A.device, dtype=out_dtype)
    ldd = _row_major_2d(out, "out")
    ep = getattr(_TLS, "ep", None)
    if ep is None:
        ep = _TLS.ep = GemmEpilogue()
    odt = out.dtype
    if odt is not torch.float32 and odt is not torch.bfloat16:
        raise VtError("GEMM output must be bf16 or fp32")
    ep.out_fp32 = 1 if odt is torch.float32 else 0
    ep.accumulate = 1 if accumulate else 0
    ep.split_k = split_k
    ep.act = act
    ep.alpha = alpha
    ep.alpha_dev = None if alpha_dev is None else alpha_dev.data_ptr()
    if bias is not None:
        assert bias.dtype is torch.float32 and bias.is_contiguous() and bias.numel() >= N
        ep.bias = bias.data_ptr()
    else:
        ep.bias = None
    if residual is not None:
        assert residual.dtype is torch.bfloat16
        ep.residual, ep.ldr = residual.data_ptr(), _row_major_2d(residual, "residual")
    else:
        ep.residual, ep.ldr = None, 0
    if aux is not None:
        assert aux.dtype is torch.bfloat16
        ep.aux, ep.ldaux = aux.data_ptr(), _row_major_2d(aux, "aux")
    else:
        ep.aux, ep.ldaux = None, 0
    if aux_out is not None:
        assert aux_out.dtype is torch.bfloat16
        ep.aux_out, ep.ldaux_out = aux_out.data_ptr(), _row_major_2d(aux_out, "aux_out")
    else:
        ep.aux_out, ep.ldaux_out = None, 0
    ep.dropout_p, ep.seed, ep.dropout_stream = dropout_p, seed, dropout_stream
    ep.remap_rows_in, ep.remap_rows_out, ep.remap_row_off = remap if remap is not None else (0, 0, 0)
    ep.remap_index = None if remap_index is None else remap_index.data_ptr()
    if colsum is not None:
        if not a_major or colsum.dtype is not torch.float32 or not colsum.is_contiguous() or colsum.numel() < M:
            raise VtError("colsum: fp32 contiguous [M] next to a [K, M] A operand")
        ep.colsum = colsum.data_ptr()
    else:
        ep.colsum = None
    if GEMM_EVENTS is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    rc = lib.vt_gemm_bf16(A.data_ptr(), lda, a_major, B.data_ptr(), ldb, b_major, out.data_ptr(), ldd, M, N, K, C.byref(ep), stream_ptr())
    if rc:
        check(rc, "vt_gemm_bf16")
    if GEMM_EVENTS is not None:
        e1.record()
        GEMM_EVENTS.append((e0, e1, M, N, K, a_major, getattr(ACTIVE_LIMIT, "v", None)))
    return out
