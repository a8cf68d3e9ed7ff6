"""GPU parity for the extension-field element layouts: fft::* with E = Quad/CubeExtension<B>
(math/src/fft/mod.rs:85,171,274,362 generic over E), ColMatrix<E>::interpolate_columns, RowMatrix<E>::evaluate_polys_over
and the auxiliary-segment commitment (prover/src/lib.rs:333-334), against the CPU oracle. Bit-exact."""
import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu
F64, F128, B3, SHA3 = 0, 1, 0, 1
CASES = [(F64, 2), (F64, 3), (F128, 2)]


@pytest.fixture(scope="module")
def wg():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


def rand_ext(field, seed, n, deg):
    """[n, deg(,2)] uniform elements of the degree-deg extension."""
    if field == F64:
        return np.ascontiguousarray(fx.synth_f64(seed, deg, n).T)
    return np.ascontiguousarray(np.moveaxis(fx.synth_f128(seed, deg, n), 0, 1))


def other_offset(field):
    return np.array([0x123456789ABCDEF % fx.P64], dtype=np.uint64) if field == F64 else np.array([12345, 678], dtype=np.uint64)


@pytest.mark.parametrize("field,deg", CASES)
def test_fft_functions_ext(wg, oracle, field, deg):
    gen = wg.matrix.GENERATOR_INTERNAL[field]
    for n, blowup, off in [(2, 2, gen), (16, 8, gen), (256, 4, other_offset(field)), (4096, 8, gen), (8192, 2, gen), (1 << 15, 8, other_offset(field))]:
        p = rand_ext(field, 300 + n + deg, n, deg)
        want = oracle.evaluate_poly_with_offset_ext(field, deg, p, off, blowup)
        got = wg.fft.evaluate_poly_with_offset(field, p, None, off, blowup, ext_degree=deg)
        assert got.shape == want.shape and np.array_equal(got, want), (n, blowup)
        # evaluate_poly (in place, no offset) == blowup 1, offset 1
        one = np.array([0xFFFFFFFF], dtype=np.uint64) if field == F64 else np.array([1, 0], dtype=np.uint64)
        e = p.copy()
        wg.fft.evaluate_poly(field, e, None, ext_degree=deg)
        assert np.array_equal(e, oracle.evaluate_poly_with_offset_ext(field, deg, p, one, 1)), n
        # interpolate_poly / interpolate_poly_with_offset (in place) bring the evaluations back to the coefficients
        wg.fft.interpolate_poly(field, e, None, ext_degree=deg)
        assert np.array_equal(e, p), n
        assert np.array_equal(e, oracle.interpolate_poly_ext(field, deg, oracle.evaluate_poly_with_offset_ext(field, deg, p, one, 1))), n
        c = wg.fft.evaluate_poly_with_offset(field, p, None, off, 1, ext_degree=deg)
        want_c = oracle.interpolate_poly_ext(field, deg, c, off)
        wg.fft.interpolate_poly_with_offset(field, c, None, off, ext_degree=deg)
        assert np.array_equal(c, want_c) and np.array_equal(c, p), n


def test_ext_degree_validation(wg):
    from starkpack_winterfell_b200._lib import ERR_UNSUPPORTED

    p = rand_ext(F64, 1, 16, 2)
    for deg in (0, 4):
        with pytest.raises(wg.WinterGpuError) as e:
            wg.fft.interpolate_poly(F64, p, None, ext_degree=deg)
        assert e.value.code == ERR_UNSUPPORTED


@pytest.mark.parametrize("field,deg", CASES)
@pytest.mark.parametrize("hash_id", [B3, SHA3])
def test_aux_segment_commitment(wg, oracle, field, deg, hash_id):
    """build_trace_commitment::<E>: equal to the base-field commitment of the coordinate columns
    (col_matrix.rs:113-119 get_base_element; row_matrix.rs:33-41)."""
    for n, ncols, blowup in [(8, 1, 8), (1 << 10, 3, 8), (1 << 13, 5, 4)]:
        cols = np.stack([rand_ext(field, 900 + 17 * c + n, n, deg) for c in range(ncols)])  # [C, n, deg(,2)]
        base = oracle.split_ext_columns(field, cols)
        want = oracle.commit_trace(field, hash_id, [base], blowup)
        cm = wg.matrix.ColMatrix(field, list(cols), ext_degree=deg)
        assert cm.num_base_cols() == ncols * deg
        ldes, tree, polys = wg.prover.build_trace_commitment(hash_id, [cm], wg.matrix.StarkDomain(field, n, blowup))
        assert np.array_equal(np.stack(polys[0].columns), oracle.merge_ext_columns(field, want["polys"][0], deg)), (n, ncols)
        assert ldes[0].elements_per_row == ncols * deg and ldes[0].row_width == (ncols * deg + 7) // 8 * 8
        assert np.array_equal(ldes[0].data, want["lde"][0]), (n, ncols)
        assert np.array_equal(tree.leaves(), want["leaves"]) and np.array_equal(tree.nodes[1:], want["nodes"][1:])
        assert tree.root() == bytes(want["root"])
        # the matrix-level functions on their own
        pm = cm.interpolate_columns()
        assert np.array_equal(np.stack(pm.columns), np.stack(polys[0].columns))
        rm = wg.matrix.RowMatrix.evaluate_polys_over(pm, wg.matrix.StarkDomain(field, n, blowup))
        assert np.array_equal(rm.data, want["lde"][0])
        assert np.array_equal(rm.commit_to_rows(hash_id).nodes[1:], want["nodes"][1:])


def test_aux_segment_aggregated(wg, oracle):
    """two aggregated extension-field segments (StarkPack shape) in one tree."""
    n, deg, blowup = 1 << 10, 2, 8
    trs = [np.stack([rand_ext(F64, 50 + 7 * t + c, n, deg) for c in range(2 + t)]) for t in range(2)]
    want = oracle.commit_trace(F64, B3, [oracle.split_ext_columns(F64, t) for t in trs], blowup)
    cms = [wg.matrix.ColMatrix(F64, list(t), ext_degree=deg) for t in trs]
    ldes, tree, _ = wg.prover.build_trace_commitment(B3, cms, wg.matrix.StarkDomain(F64, n, blowup))
    for l, w in zip(ldes, want["lde"]):
        assert np.array_equal(l.data, w)
    assert tree.root() == bytes(want["root"])
