"""argtypes / restype for every symbol declared in include/visitron_b200.h."""
import ctypes as C

P, I, I64, F, U64, U32 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64, C.c_uint32

SIGNATURES = {
    "vt_init": (I, [I]),
    "vt_num_sms": (I, []),
    "vt_launch_count": (I64, []),
    "vt_reset_launch_count": (None, []),
    "vt_gemm_bf16": (I, [P, I64, I, P, I64, I, P, I64, I64, I64, I64, P, P]),
}


def declare(lib):
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
