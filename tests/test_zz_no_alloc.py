"""vt_colsum_bf16_static: the column sums of a bf16 matrix without a caller workspace and without any allocation (static device
scratch inside the library) against vt_colsum_bf16, through the C ABI, twice in a row.  No counterpart in the reference;
the file name keeps it last in collection order."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from visitron_b200 import lib

DEV = "cuda"


def _both(L, x, ncols, out0, accumulate):
    """(existing entry point with a workspace, new entry point) on the same input, both starting from out0."""
    rows = x.size(0)
    ws = torch.empty(int(L.vt_colsum_partials_rows(rows)) * ((ncols + 7) // 8 * 8), device=DEV, dtype=torch.float32)
    ref, got = out0.clone(), out0.clone()
    torch.cuda.synchronize()
    lib.check(L.vt_colsum_bf16(x.data_ptr(), x.stride(0), rows, ncols, ws.data_ptr(), ref.data_ptr(), accumulate, lib.stream_ptr()), "vt_colsum_bf16")
    before = torch.cuda.memory_allocated()
    lib.check(L.vt_colsum_bf16_static(x.data_ptr(), x.stride(0), rows, ncols, got.data_ptr(), accumulate, lib.stream_ptr()), "vt_colsum_bf16_static")
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    return ref, got


def test_static_scratch_column_sums_match_the_workspace_entry_point_twice_in_a_row():
    L = lib.init()
    assert int(L.vt_colsum_bf16_static_max_cols()) == 4096
    torch.manual_seed(3)
    # first call: wide and tall (fills most of the scratch); second call: narrower and shorter, so that stale partials of the
    # first call lie right behind (and between) the ones it writes; third: odd column count, accumulate into a live buffer
    cases = [(20000, 3072, 3072, 0), (517, 768, 768, 0), (20000, 2304, 2304, 1), (3, 1608, 1601, 1), (9000, 4096, 4096, 0)]
    for rows, ld, ncols, accumulate in cases:
        x = (torch.randn(rows, ld, device=DEV) * 0.5 + 0.1).to(torch.bfloat16)
        out0 = torch.randn(ncols + 5, device=DEV)
        ref, got = _both(L, x, ncols, out0, accumulate)
        exact = x[:, :ncols].float().sum(0) + (out0[:ncols] if accumulate else 0)
        tol = 1e-3 + 1e-3 * exact.abs()
        assert bool(((got[:ncols] - ref[:ncols]).abs() <= tol).all()), float((got[:ncols] - ref[:ncols]).abs().max())
        assert bool(((got[:ncols] - exact).abs() <= tol + 2e-2).all())
        assert torch.equal(got[ncols:], out0[ncols:])      # nothing written past the columns asked for
        if not accumulate:
            assert torch.equal(got[:ncols], ref[:ncols])   # same kernels, same order of additions: bit-identical
    # the same call again gives the same answer (the scratch keeps the previous call's contents; none of it is read)
    x = (torch.randn(517, 768, device=DEV) * 0.5).to(torch.bfloat16)
    out0 = torch.zeros(768, device=DEV)
    a = _both(L, x, 768, out0, 0)[1]
    b = _both(L, x, 768, out0, 0)[1]
    assert torch.equal(a, b)


def test_static_scratch_rejects_what_does_not_fit():
    L = lib.init()
    x = torch.zeros(4, 4104, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros(4104, device=DEV)
    rc = L.vt_colsum_bf16_static(x.data_ptr(), x.stride(0), 4, 4104, out.data_ptr(), 0, lib.stream_ptr())
    assert rc != 0 and b"static scratch" in L.vt_last_error()
    with pytest.raises(lib.VtError):
        lib.check(L.vt_colsum_bf16_static(x.data_ptr(), x.stride(0), 4, 4097, out.data_ptr(), 0, lib.stream_ptr()), "vt_colsum_bf16_static")
    lib.check(L.vt_colsum_bf16_static(x.data_ptr(), x.stride(0), 4, 0, out.data_ptr(), 0, lib.stream_ptr()), "vt_colsum_bf16_static")   # nothing to do
