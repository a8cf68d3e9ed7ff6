"""Transforms longer than the two-pass plan (csrc/split.cuh): one more decimation step by R = 2..16 around the on-chip
transforms, recursively, up to the field's two-adicity (math/src/fft/mod.rs:291-295 accepts every power of two).

The split threshold (wg_set_split_log, default 2^24) is lowered so that every radix and the recursion run at sizes the CPU
oracle finishes in seconds; one test runs a real 2^25-point trace with the default threshold. Bit-exact against the oracle:
polys, LDE words, leaves, nodes, root."""
import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu

F64, F128 = 0, 1
B3, SHA3 = 0, 1


@pytest.fixture(scope="module")
def wg():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


@pytest.fixture
def split13(wg):
    from starkpack_winterfell_b200._lib import lib

    prev = lib().wg_set_split_log(13)
    assert prev == 24
    yield
    assert lib().wg_set_split_log(prev) == 13


def rand_cols(field, seed, cols, n):
    return fx.synth_f64(seed, cols, n) if field == F64 else fx.synth_f128(seed, cols, n)


def test_threshold_argument(wg):
    from starkpack_winterfell_b200._lib import ERR_ARG, lib

    assert lib().wg_set_split_log(0) == 24  # read only
    assert lib().wg_set_split_log(12) == ERR_ARG and lib().wg_set_split_log(25) == ERR_ARG
    assert lib().wg_set_split_log(0) == 24


# log_n - 13 = 1, 2, 3, 4 -> radix 2, 4, 8, 16; 18 and 19 split twice (16 x 2, 16 x 4)
@pytest.mark.parametrize("field,cols,log_n,blowup", [
    (F64, 3, 14, 8), (F64, 9, 15, 2), (F64, 2, 16, 4), (F64, 5, 17, 2), (F64, 2, 18, 2), (F64, 1, 19, 2),
    (F128, 2, 14, 8), (F128, 3, 15, 2), (F128, 1, 16, 4), (F128, 2, 17, 2), (F128, 1, 18, 2),
])
def test_split_interpolate_and_lde(wg, oracle, split13, field, cols, log_n, blowup):
    n = 1 << log_n
    tr = rand_cols(field, 900 + log_n + cols, cols, n)
    cm = wg.matrix.ColMatrix(field, list(tr))
    polys = cm.interpolate_columns()
    want = oracle.commit_trace(field, B3, [tr], blowup)
    for c in range(cols):
        assert np.array_equal(polys.columns[c], want["polys"][0][c]), ("poly", c)
        assert np.array_equal(cm.columns[c], tr[c])
    lde = wg.matrix.RowMatrix.evaluate_polys(polys, blowup)
    assert lde.row_width == (cols + 7) // 8 * 8
    assert np.array_equal(lde.data, want["lde"][0])  # pad columns included (zeros)


@pytest.mark.parametrize("field,h,shapes,log_n", [(F64, B3, [5, 3], 15), (F128, SHA3, [2], 14), (F64, B3, [12], 17)])
def test_split_commit_trace(wg, oracle, split13, field, h, shapes, log_n):
    """the whole commit (host entry and resident entry) with every transform on the split plan"""
    n, blowup = 1 << log_n, 8
    traces = [rand_cols(field, 950 + 7 * i + c, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, h, traces, blowup)
    cms = [wg.matrix.ColMatrix(field, list(t)) for t in traces]
    ldes, tree, polys = wg.prover.build_trace_commitment(h, cms, wg.matrix.StarkDomain(field, n, blowup))
    for t in range(len(traces)):
        for c in range(shapes[t]):
            assert np.array_equal(polys[t].columns[c], want["polys"][t][c])
        assert np.array_equal(ldes[t].data, want["lde"][t])
    assert np.array_equal(tree.nodes, want["nodes"])
    assert tree.root() == bytes(want["root"])


def test_split_offset_evaluation(wg, oracle, split13):
    """evaluate_poly_with_offset / interpolate_poly_with_offset (math/src/fft/mod.rs:150-236) through the split plan, with an
    offset that is not the generator"""
    n, blowup = 1 << 15, 4
    p = np.ascontiguousarray(rand_cols(F64, 977, 1, n)[0])
    off = np.array([(123456789 << 64) % fx.P64], dtype=np.uint64)  # Montgomery word of 123456789
    got = wg.fft.evaluate_poly_with_offset(F64, p, None, off, blowup)
    assert np.array_equal(got, oracle.evaluate_poly_with_offset(F64, p, off, blowup))
    # every blowup-th point is the coset offset * <w_n>: interpolating it over that coset gives the polynomial back
    back = np.ascontiguousarray(got[::blowup])
    wg.fft.interpolate_poly_with_offset(F64, back, None, off)
    assert np.array_equal(back, p)


def test_preferred_split_for_lde(wg, oracle):
    """2^23 / 2^24-row LDEs prefer the split plan when the scratch allows (wg_set_split_preferred_log, default 22) and fall
    back to the two-pass plan otherwise; exercised at 2^15 rows with the preference lowered to 2^13: host entry (scratch at
    the preferred size), device entry with the two-pass plan's scratch only, and the sharded commit on emulated ranks."""
    import ctypes as C

    import torch

    from starkpack_winterfell_b200 import sharded
    from starkpack_winterfell_b200._lib import check, lib

    L = lib()
    assert L.wg_set_split_preferred_log(0) == 22
    assert L.wg_set_split_preferred_log(13) == 22
    try:
        n, blowup, cols = 1 << 15, 8, 10
        tr = rand_cols(F64, 985, cols, n)
        want = oracle.commit_trace(F64, B3, [tr], blowup)
        polys = wg.matrix.ColMatrix(F64, list(tr)).interpolate_columns()
        launches0 = L.wg_launch_count()
        lde = wg.matrix.RowMatrix.evaluate_polys(polys, blowup)
        assert np.array_equal(lde.data, want["lde"][0])
        assert L.wg_launch_count() - launches0 == 4  # de-interleave, two passes over 4 x 10 sub-columns, combine
        # device entry, scratch of the two-pass plan only: same result through the two passes
        dev = torch.device("cuda", 0)
        d_polys = torch.from_numpy(np.ascontiguousarray(want["polys"][0]).view(np.int64)).to(dev)
        rw = 16
        d_lde = torch.empty((n * blowup, rw), dtype=torch.int64, device=dev)
        se = cols * n * blowup
        assert L.wg_lde_scratch_elems(F64, cols, n, blowup) == cols * n * (2 * blowup + 1)
        scr = torch.empty(se, dtype=torch.int64, device=dev)
        off = np.array([(7 << 64) % fx.P64], dtype=np.uint64)
        launches0 = L.wg_launch_count()
        check(L.wg_lde_dev(F64, C.c_void_p(d_polys.data_ptr()), cols, n, blowup, C.c_void_p(off.ctypes.data),
                           C.c_void_p(d_lde.data_ptr()), rw, C.c_void_p(scr.data_ptr()), se, None))
        torch.cuda.synchronize()
        assert L.wg_launch_count() - launches0 == 2
        assert np.array_equal(d_lde.cpu().numpy().view(np.uint64), want["lde"][0])
        # sharded commit on emulated ranks (chunked forms take the split plan, the scatter form the two-pass plan)
        d_cols = torch.from_numpy(np.ascontiguousarray(tr).view(np.int64)[..., None]).to(dev)
        for exchange, chunks in ((1, 0), (2, 2), (3, 0)):
            root, leaves, nodes = sharded.commit_emulated(F64, B3, n, blowup, cols, 2, d_cols, exchange=exchange, chunks=chunks)
            assert root == bytes(want["root"]) and np.array_equal(nodes, want["nodes"])
    finally:
        assert L.wg_set_split_preferred_log(22) == 13


def test_beyond_2_24_default_threshold(wg, oracle):
    """a real 2^25-row column (the two-pass plan stops at 2^24): radix-2 step around two 2^24-point transforms"""
    n, blowup = 1 << 25, 2
    tr = rand_cols(F64, 991, 1, n)
    cm = wg.matrix.ColMatrix(F64, list(tr))
    polys = cm.interpolate_columns()
    want = oracle.commit_trace(F64, B3, [tr], blowup, want=("polys", "lde", "root"))
    assert np.array_equal(polys.columns[0], want["polys"][0][0])
    lde = wg.matrix.RowMatrix.evaluate_polys(polys, blowup)
    assert np.array_equal(lde.data, want["lde"][0])
    leaves = lde.hash_rows(B3)
    tree = wg.merkle.MerkleTree.new(B3, leaves)
    assert tree.root() == bytes(want["root"])
