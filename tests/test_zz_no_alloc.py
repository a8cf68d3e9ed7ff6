"""wg_interpolate_columns_noalloc_dev: the no-allocation variant of wg_interpolate_columns_dev (static __device__ scratch,
include/winter_gpu.h). No counterpart in the reference. Called through the
C ABI and compared bit for bit (integer field arithmetic: tolerance zero, as in every other parity test here) with the
existing entry point, on two calls in a row with different inputs (the static scratch keeps the previous call's
contents), plus the size limit and the not-resident-tables error. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from tests import fixtures as fx

pytestmark = pytest.mark.gpu


def _dev(a, field):
    a = np.ascontiguousarray(a).view(np.int64)
    return torch.from_numpy(a[..., None] if field == 0 else a).to("cuda:0")


def _existing(L, field, cols, n):
    polys = torch.empty_like(cols)
    se = int(L.wg_interpolate_scratch_elems(field, cols.shape[0], n))
    scratch = torch.empty((max(se, 1), cols.shape[2]), dtype=torch.int64, device=cols.device)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.wg_interpolate_columns_dev(field, C.c_void_p(cols.data_ptr()), cols.shape[0], n, C.c_void_p(polys.data_ptr()),
                                      C.c_void_p(scratch.data_ptr()), se, s)
    assert rc == 0, L.wg_last_error()
    return polys


def _noalloc(L, field, cols, n, polys):
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return L.wg_interpolate_columns_noalloc_dev(field, C.c_void_p(cols.data_ptr()), cols.shape[0], n, C.c_void_p(polys.data_ptr()), s)


@pytest.mark.parametrize("field,n_cols,n", [(0, 8, 1 << 18), (0, 5, 1 << 13), (1, 3, 1 << 15), (0, 7, 1024)])
def test_noalloc_matches_existing_entry_point_twice(field, n_cols, n):
    from starkpack_winterfell_b200 import _lib

    _lib.init(0)
    L = _lib.lib()
    outs = []
    for call in range(2):  # two calls in a row, different data: the second must not see the first one's scratch contents
        tr = fx.synth_f64(900 + call, n_cols, n) if field == 0 else fx.synth_f128(900 + call, n_cols, n)
        cols = _dev(tr, field)
        want = _existing(L, field, cols, n)  # also makes the twiddle tables of this length resident
        got = torch.full_like(cols, -1)  # allocated BEFORE the call under test; the call itself allocates nothing
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        free0, _ = torch.cuda.mem_get_info()
        rc = _noalloc(L, field, cols, n, got)
        torch.cuda.synchronize()
        assert rc == 0, L.wg_last_error()
        free1, _ = torch.cuda.mem_get_info()
        assert torch.cuda.memory_allocated() == before
        if call == 1:  # (the first call may make the driver load the module that holds the static array: not an allocation call)
            assert free1 >= free0, "the no-allocation entry point allocated device memory"
        assert torch.equal(got, want), (field, n, call)
        outs.append(got)
    assert not torch.equal(outs[0], outs[1])


def test_noalloc_rejects_oversize_and_cold_tables():
    from starkpack_winterfell_b200 import _lib

    _lib.init(0)
    L = _lib.lib()
    n = 1 << 18
    cols = torch.zeros((9, n, 1), dtype=torch.int64, device="cuda:0")  # 9 * 2^18 * 8 B = 18 MiB > 16 MiB
    out = torch.empty_like(cols)
    assert _noalloc(L, 0, cols, n, out) == _lib.ERR_TOO_LARGE
    assert b"static scratch" in L.wg_last_error()
    # a length whose tables nobody has built in this process: refused instead of allocating them
    n2 = 1 << 17
    c2 = torch.zeros((1, n2, 2), dtype=torch.int64, device="cuda:0")
    rc = _noalloc(L, 1, c2, n2, torch.empty_like(c2))
    if rc != 0:  # (an earlier test of the same process may already have used this f128 length)
        assert rc == _lib.ERR_ARG and b"not resident" in L.wg_last_error()
    assert L.wg_interpolate_columns_noalloc_dev(0, None, 1, n, C.c_void_p(out.data_ptr()), None) == _lib.ERR_ARG  # null input
