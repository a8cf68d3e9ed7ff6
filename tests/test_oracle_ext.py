"""CPU pinning of the extension-field restatement (oracle.c or_*_ext): the reference's FFT on E = Quad/CubeExtension<B>
multiplies by base-field twiddles only (math/src/fft/fft_inputs.rs:116-123 `mul_base`, extensions/quadratic.rs:187-190,
cubic.rs:196-199: coordinate-wise), so its result is the base-field definition applied to every coordinate. Checked
against the big-int O(n^2) definitions (polynom::eval_many, math/src/fft/tests.rs:19-60 style)."""
import numpy as np
import pytest

from oracle import pyref

F64, F128 = 0, 1


def rand_ext(field, seed, n, deg):
    p = pyref.MOD[field]
    rng = np.random.default_rng(seed)
    return [[int.from_bytes(rng.bytes(16), "little") % p for _ in range(deg)] for _ in range(n)]


def to_arr(field, rows):
    n, deg = len(rows), len(rows[0])
    flat = pyref.ints_to_array(field, [v for r in rows for v in r])
    return flat.reshape((n, deg) if field == F64 else (n, deg, 2))


def to_ints(field, arr):
    deg = arr.shape[1]
    flat = pyref.array_to_ints(field, arr)
    return [flat[i * deg:(i + 1) * deg] for i in range(arr.shape[0])]


@pytest.mark.parametrize("field,deg", [(F64, 2), (F64, 3), (F128, 2)])
def test_ext_evaluate_and_interpolate(oracle, field, deg):
    for n, blowup in ((8, 4), (32, 2), (64, 8)):
        coeffs = rand_ext(field, 5 * n + deg, n, deg)
        off = pyref.GENERATOR[field]
        dom = pyref.lde_domain(field, n, blowup, off)
        want = list(zip(*[pyref.eval_many(field, [c[d] for c in coeffs], dom) for d in range(deg)]))
        off_arr = pyref.ints_to_array(field, [off]).reshape(-1)
        got = oracle.evaluate_poly_with_offset_ext(field, deg, to_arr(field, coeffs), off_arr, blowup)
        assert to_ints(field, got) == [list(r) for r in want], (n, blowup)
        # interpolate back: evaluations over the trace domain (offset 1) and over the coset
        ev1 = oracle.evaluate_poly_with_offset_ext(field, deg, to_arr(field, coeffs), pyref.ints_to_array(field, [1]).reshape(-1), 1)
        assert to_ints(field, oracle.interpolate_poly_ext(field, deg, ev1)) == coeffs
        evc = oracle.evaluate_poly_with_offset_ext(field, deg, to_arr(field, coeffs), off_arr, 1)
        assert to_ints(field, oracle.interpolate_poly_ext(field, deg, evc, off_arr)) == coeffs


def test_split_merge_columns(oracle):
    a = np.arange(2 * 8 * 3, dtype=np.uint64).reshape(2, 8, 3)
    b = oracle.split_ext_columns(F64, a)
    assert b.shape == (6, 8) and list(b[4]) == [int(a[1, r, 1]) for r in range(8)]
    assert np.array_equal(oracle.merge_ext_columns(F64, b, 3), a)
    c = np.arange(2 * 4 * 2 * 2, dtype=np.uint64).reshape(2, 4, 2, 2)
    d = oracle.split_ext_columns(F128, c)
    assert d.shape == (4, 4, 2) and np.array_equal(d[3], c[1, :, 1, :])
    assert np.array_equal(oracle.merge_ext_columns(F128, d, 2), c)
