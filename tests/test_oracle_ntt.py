"""Pins oracle.c NTT / LDE / Merkle / commit against the Python big-int definitions the reference's own
tests use (O(n^2) eval_many): math/src/fft/tests.rs:19-72, doc-tests fft/mod.rs:147-170,253-273,
prover/src/matrix/tests.rs:14-40, prover/src/trace/tests.rs:42-128, crypto/src/merkle/tests.rs:68-99."""
import random

import numpy as np
import pytest

from oracle import pyref
from tests import fixtures as fx


def rand_ints(field, n, seed):
    rnd = random.Random(seed)
    return [rnd.randrange(pyref.MOD[field]) for _ in range(n)]


@pytest.mark.parametrize("field", [0, 1])
@pytest.mark.parametrize("n", [2, 4, 8, 16, 64, 512, 2048])
def test_twiddles(oracle, field, n):
    for inv in (False, True):
        got = pyref.array_to_ints(field, oracle.get_twiddles(field, n, inv))
        assert got == pyref.twiddles(field, n, inv)


@pytest.mark.parametrize("field", [0, 1])
@pytest.mark.parametrize("n", [2, 4, 8, 16, 128, 512])
def test_interpolate_and_evaluate(oracle, field, n):
    p = pyref.MOD[field]
    coeffs = rand_ints(field, n, n + field)
    w = pyref.root_of_unity(field, n.bit_length() - 1)
    evals = pyref.eval_many(field, coeffs, [pow(w, k, p) for k in range(n)])
    got = oracle.interpolate_poly(field, pyref.ints_to_array(field, evals))
    assert pyref.array_to_ints(field, got) == coeffs
    for blowup in (1, 2, 8):
        if n * blowup > 1024:
            continue
        off = pyref.GENERATOR[field]
        out = oracle.evaluate_poly_with_offset(field, pyref.ints_to_array(field, coeffs),
                                               pyref.ints_to_array(field, [off])[0], blowup)
        assert pyref.array_to_ints(field, out) == pyref.eval_many(field, coeffs, pyref.lde_domain(field, n, blowup))


def test_merkle_fixed_leaves(oracle):
    for leaves in (fx.LEAVES4, fx.LEAVES8):
        lb = [bytes(l) for l in leaves]
        for h in (0, 1):
            nodes = oracle.merkle_nodes(h, np.array(leaves, dtype=np.uint8))
            ref = pyref.merkle_nodes(h, lb)
            assert [bytes(x) for x in nodes] == ref
            assert bytes(nodes[0]) == bytes(32)
    # nested-merge root, as crypto/src/merkle/tests.rs:68-80
    h2 = lambda a, b: pyref.hash_bytes(0, a + b)
    lb = [bytes(l) for l in fx.LEAVES4]
    assert bytes(oracle.merkle_nodes(0, np.array(fx.LEAVES4, dtype=np.uint8))[1]) == h2(h2(lb[0], lb[1]), h2(lb[2], lb[3]))
    with pytest.raises(ValueError):
        oracle.merkle_nodes(0, np.zeros((1, 32), dtype=np.uint8))
    with pytest.raises(ValueError):
        oracle.merkle_nodes(0, np.zeros((6, 32), dtype=np.uint8))


def _check_commit(oracle, field, hash_id, traces_int, blowup):
    arrs = [np.stack([pyref.ints_to_array(field, col) for col in tr]) for tr in traces_int]
    got = oracle.commit_trace(field, hash_id, arrs, blowup)
    polys, ldes, leaves, nodes = pyref.commit(field, hash_id, traces_int, blowup)
    for t, tr in enumerate(traces_int):
        C = len(tr)
        for c in range(C):
            assert pyref.array_to_ints(field, got["polys"][t][c]) == polys[t][c]
        lde = got["lde"][t]
        assert lde.shape[1] == (C + 7) // 8 * 8
        for c in range(C):
            assert pyref.array_to_ints(field, lde[:, c]) == ldes[t][c]
        assert not lde[:, C:].any()  # F7: zero padding
    assert [bytes(x) for x in got["leaves"]] == leaves
    assert [bytes(x) for x in got["nodes"]] == nodes
    assert bytes(got["root"]) == nodes[1]
    return got


def test_fib_trace_commit(oracle):
    """prover/src/trace/tests.rs:42-128: 8x2 fib trace, blowup 8, Blake3_256."""
    got = _check_commit(oracle, 1, 0, [fx.FIB_COLS], 8)
    # polys evaluate back to the trace over the trace domain (tests.rs:42-60)
    w = pyref.root_of_unity(1, 3)
    for c in range(2):
        poly = pyref.array_to_ints(1, got["polys"][0][c])
        assert pyref.eval_many(1, poly, [pow(w, k, pyref.P128) for k in range(8)]) == fx.FIB_COLS[c]


@pytest.mark.parametrize("field,hash_id,cols,n,blowup", [
    (0, 0, 20, 64, 8),   # prover/src/matrix/tests.rs:14-40 shape, scaled (f64, ragged segment: 20 = 2*8+4)
    (0, 1, 3, 256, 2),
    (0, 0, 8, 16, 4),
    (1, 1, 4, 32, 8),    # rescue-like: f128, Sha3_256
    (1, 0, 10, 16, 2),   # do_work-like: 10 f128 columns
])
def test_commit_vs_bigint(oracle, field, hash_id, cols, n, blowup):
    tr = [rand_ints(field, n, 100 * c + n) for c in range(cols)]
    _check_commit(oracle, field, hash_id, [tr], blowup)


def test_aggregated_commit(oracle):
    """fork-specific commit_to_comb_rows (row_matrix.rs:204-238): one tree over T traces' concatenated rows."""
    traces = [[rand_ints(1, 16, 7 * t + c) for c in range(2)] for t in range(3)]
    _check_commit(oracle, 1, 0, traces, 4)
    traces = [[rand_ints(0, 32, 11 * t + c) for c in range(3 + t)] for t in range(2)]
    _check_commit(oracle, 0, 0, traces, 2)


@pytest.mark.parametrize("field", [0, 1])
@pytest.mark.parametrize("log_n", [10, 11])
def test_concurrent_segment_path_equals_serial_columns(oracle, field, log_n):
    """The oracle's matrix LDE restates the reference's `concurrent` segment transform (sqrt(n) decomposition with nested
    parallel row FFTs, prover/src/matrix/segments.rs:232-298) from n = 1024 up; every column of its result must equal the
    serial fft::evaluate_poly_with_offset of that column (math/src/fft/serial.rs:26-57), for an even and an odd log2 n
    (square and 1x2-stretched transposes). prover/src/matrix/tests.rs:14-40 checks the same property."""
    n, blowup, cols = 1 << log_n, 4, 11
    tr = fx.synth_f64(77 + log_n, cols, n) if field == 0 else fx.synth_f128(77 + log_n, cols, n)
    r = oracle.commit_trace(field, 0, [tr], blowup, want=("polys", "lde"))
    off = pyref.ints_to_array(field, [pyref.GENERATOR[field]])[0]
    for c in (0, 5, 10):
        want = oracle.evaluate_poly_with_offset(field, r["polys"][0][c], off, blowup)
        assert np.array_equal(r["lde"][0][:, c], want)
    assert not r["lde"][0][:, cols:].any()
