"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/visitron_b200.h declares (no compute calls -- there is no GPU here), and the product path
fails loudly without a device instead of falling back."""
import os
import re

import pytest
import torch

from visitron_b200 import _signatures, build, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _prototypes():
    hdr = open(os.path.join(ROOT, "include", "visitron_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return re.findall(r"\b(?:int|int64_t|void|const char\*)\s+(vt_\w+)\s*\(([^;]*?)\)\s*;", hdr, flags=re.S)


def test_library_builds_and_exports_every_declared_symbol():
    if not os.path.exists(lib.LIB_PATH):
        build.build(verbose=False)
    L = lib.load()
    protos = _prototypes()
    assert len(protos) >= 30
    for name, args in protos:
        assert hasattr(L, name), "symbol %s declared in the header is not exported" % name
        n = 0 if args.strip() in ("void", "") else len(args.split(","))
        if name != "vt_last_error":
            assert name in _signatures.SIGNATURES, name
            assert len(_signatures.SIGNATURES[name][1]) == n, name


def test_epilogue_struct_layout_matches_header():
    # field order of vt_gemm_epilogue in the header == ctypes mirror
    hdr = open(os.path.join(ROOT, "include", "visitron_b200.h")).read()
    body = hdr[hdr.index("typedef struct {"):hdr.index("} vt_gemm_epilogue;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct {", "").strip()
        if not decl:
            continue
        parts = decl.split(",")
        names.append(parts[0].split()[-1].lstrip("*"))
        names += [p.strip().lstrip("*") for p in parts[1:]]
    assert names == [f[0] for f in lib.GemmEpilogue._fields_]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_no_cpu_fallback():
    with pytest.raises(lib.VtError):
        lib.init()
    rc = lib.load().vt_init(0)
    assert rc != 0 and b"no CUDA device" in lib.load().vt_last_error()


def test_agent_heads_refuse_cpu_tensors():
    from visitron_b200 import agent_heads as ah
    with pytest.raises(lib.VtError):
        ah.small_linear(torch.randn(2, 8), torch.randn(4, 8))
    with pytest.raises(lib.VtError):
        ah.SoftDotAttention(8)(torch.randn(2, 8), torch.randn(2, 3, 8))
    with pytest.raises(lib.VtError):
        ah.masked_cross_entropy(torch.randn(2, 3), torch.zeros(2, dtype=torch.long))
