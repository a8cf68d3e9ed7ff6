"""GPU parity: every C-ABI entry point against the CPU oracle (oracle/, restated from the reference) on the
same seeded inputs. Bit-exact: LDE words (Montgomery form for f64), leaf digests, tree nodes, root.
Shapes follow the reference's own tests (math/src/fft/tests.rs:19-72, prover/src/matrix/tests.rs:14-40,
prover/src/trace/tests.rs:42-128, crypto/src/merkle/tests.rs:12-99)."""
import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu

F64, F128 = 0, 1
B3, SHA3 = 0, 1
P64 = fx.P64


@pytest.fixture(scope="module")
def wg():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


def to_mont(v):
    """canonical uint64 array -> Montgomery words (python ints; test-side helper)."""
    flat = [(int(x) << 64) % P64 for x in np.asarray(v).reshape(-1)]
    return np.array(flat, dtype=np.uint64).reshape(np.asarray(v).shape)


def rand_cols(field, seed, cols, n):
    # any word in [0, p) is a valid internal (Montgomery) representation, so uniform words are uniform elements
    return fx.synth_f64(seed, cols, n) if field == F64 else fx.synth_f128(seed, cols, n)


def _field_op(wg, field, op, a, b):
    import ctypes as C

    from starkpack_winterfell_b200._lib import check, lib

    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = np.ascontiguousarray(b, dtype=np.uint64)
    out = np.empty_like(a)
    check(lib().wg_debug_field_op(field, op, C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(out.ctypes.data), a.shape[0]))
    return out


def test_device_field_ops_f64(wg):
    """the PTX carry-chain add/sub/Montgomery product the kernels run, against Python big ints
    (math/src/field/f64/tests.rs:17-250 style: edge values + random)."""
    rng = np.random.default_rng(11)
    edge = [0, 1, 2, P64 - 1, P64 - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P64 - 2**32, 0xFFFFFFFF00000000]
    a = np.array(edge * len(edge) + [int(x) % P64 for x in rng.integers(0, 2**64, 4000, dtype=np.uint64)], dtype=np.uint64)
    b = np.array([e for e in edge for _ in edge] + [int(x) % P64 for x in rng.integers(0, 2**64, 4000, dtype=np.uint64)], dtype=np.uint64)
    rinv = pow(2**64, -1, P64)
    A, Bv = [int(x) for x in a], [int(x) for x in b]
    assert [int(x) for x in _field_op(wg, F64, 0, a, b)] == [(x + y) % P64 for x, y in zip(A, Bv)]
    assert [int(x) for x in _field_op(wg, F64, 1, a, b)] == [(x - y) % P64 for x, y in zip(A, Bv)]
    assert [int(x) for x in _field_op(wg, F64, 2, a, b)] == [x * y * rinv % P64 for x, y in zip(A, Bv)]
    for i in range(16):
        assert [int(x) for x in _field_op(wg, F64, 16 + i, a, b)] == [x * pow(2, 12 * i, P64) % P64 for x in A], i
        assert [int(x) for x in _field_op(wg, F64, 32 + i, a, b)] == [x * pow(2, -12 * i, P64) % P64 for x in A], i
        # the butterfly form the NTT kernels use: (a - b) * w16^(+-i), sign of the shift constants absorbed by the subtraction
        assert [int(x) for x in _field_op(wg, F64, 64 + i, a, b)] == [(x - y) * pow(2, 12 * i, P64) % P64 for x, y in zip(A, Bv)], i
        assert [int(x) for x in _field_op(wg, F64, 80 + i, a, b)] == [(x - y) * pow(2, -12 * i, P64) % P64 for x, y in zip(A, Bv)], i
    w = rng.integers(0, 2**64, 4000, dtype=np.uint64)
    assert [int(x) for x in _field_op(wg, F64, 48, w, w)] == [int(x) * rinv % P64 for x in w]


def test_device_lazy_goldilocks_ops(wg):
    """field_lazy.cuh: the lazily reduced add / sub / canon and the products with a non-canonical first operand that the
    decimation-in-time register network runs on. First operand ANY u64 (including [p, 2^64)), second operand <= p; results
    are compared modulo p, `canon` and the products exactly (they must come out canonical). The two 16-point networks (the
    canonical DIF one and the lazy DIT one) must agree element for element."""
    rng = np.random.default_rng(21)
    M = 2**64
    edge_a = [0, 1, 2, P64 - 1, P64, P64 + 1, M - 1, M - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P64 - 2**32, 0xFFFFFFFF00000000,
              0xFFFFFFFF00000001, 0xFFFFFFFFFFFFFFFE, 0xFFFFFFFF, 0x100000000, M - 2**32, M - 2**32 + 1]
    edge_b = [0, 1, 2, P64 - 1, P64, P64 - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P64 - 2**32, 0xFFFFFFFF00000000, 0xFFFFFFFF]
    ra = [int(x) for x in rng.integers(0, 2**64, 6000, dtype=np.uint64)]
    rb = [int(x) % (P64 + 1) for x in rng.integers(0, 2**64, 6000, dtype=np.uint64)]
    A = [x for x in edge_a for _ in edge_b] + ra
    Bv = [y for _ in edge_a for y in edge_b] + rb
    a, b = np.array(A, dtype=np.uint64), np.array(Bv, dtype=np.uint64)
    rinv = pow(M, -1, P64)
    for op in (96, 97):
        assert [int(x) % P64 for x in _field_op(wg, F64, op, a, b)] == [(x + y) % P64 for x, y in zip(A, Bv)], op
    assert [int(x) % P64 for x in _field_op(wg, F64, 98, a, b)] == [(x - y) % P64 for x, y in zip(A, Bv)]
    assert [int(x) for x in _field_op(wg, F64, 99, a, b)] == [x % P64 for x in A]
    # products: second operand canonical (< p)
    Bc = [y % P64 for y in Bv]
    bc = np.array(Bc, dtype=np.uint64)
    for op in (100, 103):
        assert [int(x) for x in _field_op(wg, F64, op, a, bc)] == [x * y * rinv % P64 for x, y in zip(A, Bc)], op
    assert [int(x) for x in _field_op(wg, F64, 101, a, b)] == [x * 2**28 * rinv % P64 for x in A]
    assert [int(x) for x in _field_op(wg, F64, 102, a, b)] == [x * 2**52 * rinv % P64 for x in A]
    # 16-point networks: inputs canonical; thread i returns output (i & 15)
    n = 16 * 400
    x = np.array([int(v) % P64 for v in rng.integers(0, 2**64, n, dtype=np.uint64)], dtype=np.uint64)
    y = np.array([int(v) % P64 for v in rng.integers(0, 2**64, n, dtype=np.uint64)], dtype=np.uint64)
    x[:32] = [P64 - 1, 0, 1, P64 - 2] * 8
    y[:32] = [P64 - 1] * 16 + [0] * 16
    assert np.array_equal(_field_op(wg, F64, 112, x, y), _field_op(wg, F64, 114, x, y))
    assert np.array_equal(_field_op(wg, F64, 113, x, y), _field_op(wg, F64, 115, x, y))
    # and against the definition, X[k] = sum_j v_j w16^(jk) with w16 = 2^12
    G = 0x9E3779B97F4A7C15
    got = [int(v) for v in _field_op(wg, F64, 114, x[:64], y[:64])]
    for i in range(64):
        v = [((int(y[i]) if k & 1 else int(x[i])) + (k >> 1) * G) % M for k in range(16)]
        v = [t - P64 if t >= P64 else t for t in v]
        k = i & 15
        assert got[i] == sum(v[j] * pow(2, 12 * j * k, P64) for j in range(16)) % P64


def test_device_field_ops_f128(wg):
    rng = np.random.default_rng(12)
    q = fx.P128
    K = 45 * 2**40 - 1  # 2^128 mod q
    edge = [0, 1, 2, q - 1, q - 2, 2**64 - 1, 2**64, 2**127, q - 2**64, 2**128 - 2**64, K, K + 1, q - K, 2**96, 2**96 - 1,
            q - 2**40, (q - 1) // 2, 2**32 - 1, 2**32, (2**32 - 1) * (2**96 + 2**64 + 2**32 + 1) % q]
    # every edge value meets every edge value (the reversed copy pairs vals[i] with vals[-1-i])
    vals = [e for e in edge for _ in edge] + [int.from_bytes(rng.bytes(16), "little") % q for _ in range(3000)] + [e for _ in edge for e in edge][::-1]
    a = np.array([(v & fx.MASK, v >> 64) for v in vals], dtype=np.uint64)
    bv = vals[::-1]
    b = np.array([(v & fx.MASK, v >> 64) for v in bv], dtype=np.uint64)
    un = lambda r: [int(x[0]) | (int(x[1]) << 64) for x in r]
    assert un(_field_op(wg, F128, 0, a, b)) == [(x + y) % q for x, y in zip(vals, bv)]
    assert un(_field_op(wg, F128, 1, a, b)) == [(x - y) % q for x, y in zip(vals, bv)]
    assert un(_field_op(wg, F128, 2, a, b)) == [(x * y) % q for x, y in zip(vals, bv)]
    w16 = pow(23953097886125630542083529559205016746, 1 << 36, q)
    for i in range(16):
        assert un(_field_op(wg, F128, 16 + i, a, b)) == [x * pow(w16, i, q) % q for x in vals], i
        assert un(_field_op(wg, F128, 32 + i, a, b)) == [x * pow(w16, -i, q) % q for x in vals], i
        assert un(_field_op(wg, F128, 64 + i, a, b)) == [(x - y) * pow(w16, i, q) % q for x, y in zip(vals, bv)], i
        assert un(_field_op(wg, F128, 80 + i, a, b)) == [(x - y) * pow(w16, -i, q) % q for x, y in zip(vals, bv)], i


SIZES = [2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 1 << 14, 1 << 15, 1 << 17]


@pytest.mark.parametrize("field", [F64, F128])
def test_twiddles(wg, oracle, field):
    for n in [2, 4, 8, 64, 2048, 1 << 13, 1 << 16]:
        assert np.array_equal(wg.fft.get_twiddles(field, n), oracle.get_twiddles(field, n)), n
        assert np.array_equal(wg.fft.get_inv_twiddles(field, n), oracle.get_twiddles(field, n, True)), n


@pytest.mark.parametrize("field", [F64, F128])
def test_interpolate_poly(wg, oracle, field):
    for n in SIZES:
        v = rand_cols(field, 100 + n, 1, n)[0]
        want = oracle.interpolate_poly(field, v)
        got = np.ascontiguousarray(v).copy()
        wg.fft.interpolate_poly(field, got, wg.fft.get_inv_twiddles(field, n) if n <= 64 else None)
        assert np.array_equal(got, want), n


@pytest.mark.parametrize("field", [F64, F128])
def test_evaluate_poly_with_offset(wg, oracle, field):
    gen = wg.matrix.GENERATOR_INTERNAL[field]
    other = np.array([0x123456789ABCDEF % P64], dtype=np.uint64) if field == F64 else np.array([12345, 678], dtype=np.uint64)
    for n, blowup, off in [(2, 2, gen), (8, 8, gen), (16, 4, other), (256, 8, gen), (2048, 2, other), (4096, 8, gen),
                           (8192, 8, gen), (1 << 14, 16, other), (1 << 16, 2, gen)]:
        p = rand_cols(field, 200 + n, 1, n)[0]
        want = oracle.evaluate_poly_with_offset(field, p, off, blowup)
        got = wg.fft.evaluate_poly_with_offset(field, p, None, off, blowup)
        assert np.array_equal(got, want), (n, blowup)


def _one(field):
    return np.array([0xFFFFFFFF], dtype=np.uint64) if field == F64 else np.array([1, 0], dtype=np.uint64)


@pytest.mark.parametrize("field", [F64, F128])
def test_evaluate_poly_and_serial_fft(wg, oracle, field):
    """fft::evaluate_poly / serial_fft == evaluation over the unshifted domain (math/src/fft/tests.rs:19-46)."""
    for n in [2, 4, 32, 1024, 4096, 8192, 1 << 15]:
        p = rand_cols(field, 300 + n, 1, n)[0]
        want = oracle.evaluate_poly_with_offset(field, p, _one(field), 1)
        got = np.ascontiguousarray(p).copy()
        wg.fft.evaluate_poly(field, got, wg.fft.get_twiddles(field, n) if n <= 32 else None)
        assert np.array_equal(got, want), n
        got2 = np.ascontiguousarray(p).copy()
        wg.fft.serial_fft(field, got2)
        assert np.array_equal(got2, want), n


@pytest.mark.parametrize("field", [F64, F128])
def test_interpolate_poly_with_offset_and_infer_degree(wg, oracle, field):
    """interpolate_poly_with_offset inverts evaluation over the coset (math/src/fft/tests.rs:48-60); infer_degree finds
    the highest non-zero coefficient (fft/mod.rs:524-560)."""
    gen = wg.matrix.GENERATOR_INTERNAL[field]
    other = np.array([0x123456789ABCDEF % P64], dtype=np.uint64) if field == F64 else np.array([12345, 678], dtype=np.uint64)
    for n, off in [(2, gen), (16, other), (1024, gen), (4096, other), (8192, gen), (1 << 14, other)]:
        p = rand_cols(field, 400 + n, 1, n)[0]
        evals = oracle.evaluate_poly_with_offset(field, p, off, 1)
        got = np.ascontiguousarray(evals).copy()
        wg.fft.interpolate_poly_with_offset(field, got, None, off)
        assert np.array_equal(got, p), n
        deg = max(n // 4 - 1, 0)
        q = np.ascontiguousarray(p).copy()
        q[deg + 1:] = 0
        if not q[deg].any():
            q[deg] = _one(field)
        ev = oracle.evaluate_poly_with_offset(field, q, off, 1)
        assert wg.fft.infer_degree(field, ev, off) == deg, n
    with pytest.raises(wg.WinterGpuError) as e:
        wg.fft.interpolate_poly_with_offset(F64, np.zeros(8, dtype=np.uint64), None, np.zeros(1, dtype=np.uint64))
    assert e.value.code == wg._lib.ERR_ZERO_OFFSET


@pytest.mark.parametrize("field,cols,n,blowup", [
    (F64, 64, 256, 8),      # prover/src/matrix/tests.rs:14-40
    (F64, 1, 8, 2), (F64, 7, 16, 4), (F64, 9, 64, 8), (F64, 13, 4096, 8), (F64, 100, 1024, 8),
    (F64, 10, 8192, 8), (F64, 21, 1 << 14, 4), (F64, 3, 1 << 16, 8),
    (F128, 2, 8, 8),        # prover/src/trace/tests.rs fib shape
    (F128, 4, 1024, 8), (F128, 10, 4096, 2), (F128, 5, 8192, 8), (F128, 2, 1 << 15, 8),
])
def test_interpolate_columns_and_lde(wg, oracle, field, cols, n, blowup):
    tr = rand_cols(field, 300 + cols * 7 + n, cols, n)
    cm = wg.matrix.ColMatrix(field, list(tr))
    polys = cm.interpolate_columns()
    want = oracle.commit_trace(field, B3, [tr], blowup)
    for c in range(cols):
        assert np.array_equal(polys.columns[c], want["polys"][0][c]), ("poly", c)
        assert np.array_equal(cm.columns[c], tr[c])  # input untouched (col_matrix.rs:196-206)
    lde = wg.matrix.RowMatrix.evaluate_polys(polys, blowup)
    assert lde.row_width == (cols + 7) // 8 * 8 and lde.elements_per_row == cols
    assert np.array_equal(lde.data, want["lde"][0])
    # leaves and tree over the same LDE
    for h in (B3, SHA3):
        leaves = lde.hash_rows(h)
        assert np.array_equal(leaves, oracle.hash_rows(field, h, want["lde"][0], n * blowup, lde.row_width, cols))


@pytest.mark.parametrize("field", [F64, F128])
def test_hash_rows_of_empty_rows(wg, field):
    """hash_elements(&[]) is the digest of the empty input (crypto/src/hash/blake/mod.rs:46-59, sha/mod.rs:43-56); a FRI
    remainder shorter than the blowup commits to exactly that (fri/src/prover/mod.rs:213-223)."""
    import hashlib

    import blake3

    w = 1 if field == F64 else 2
    m = np.zeros((5, 8 * w), dtype=np.uint64)
    for h, want in ((B3, blake3.blake3(b"").digest()), (SHA3, hashlib.sha3_256(b"").digest())):
        got = wg.matrix.RowMatrix(field, m.reshape(5, 8, w)[..., 0] if field == F64 else m.reshape(5, 8, 2), 8, 0).hash_rows(h)
        assert all(bytes(r) == want for r in got), h


@pytest.mark.parametrize("field,cols", [(F64, 1), (F64, 8), (F64, 16), (F64, 17), (F64, 100), (F64, 128), (F64, 129),
                                         (F64, 256), (F64, 300), (F64, 700), (F128, 1), (F128, 2), (F128, 4), (F128, 9),
                                         (F128, 64), (F128, 65), (F128, 150)])
def test_hash_rows_block_and_chunk_shapes(wg, oracle, field, cols):
    """single block, exact block, tail block, exact chunk (1024 B), 2-chunk, 3+-chunk BLAKE3 trees; SHA3 rate edges."""
    rows = 77  # ragged: not a multiple of the 32 rows a warp owns
    rw = (cols + 7) // 8 * 8
    m = rand_cols(field, 400 + cols, rw, rows)  # [rw, rows(,2)] -> row-major [rows, rw(,2)]
    m = np.ascontiguousarray(np.swapaxes(m, 0, 1))
    if field == F64:
        m[3, 0] = np.uint64(P64 - 1)  # extreme words
        m[4, 0] = np.uint64(0)
    for h in (B3, SHA3):
        want = oracle.hash_rows(field, h, m, rows, rw, cols)
        got = wg.matrix.RowMatrix(field, m, rw, cols).hash_rows(h)
        assert np.array_equal(got, want), (cols, h)


@pytest.mark.parametrize("h", [B3, SHA3])
def test_merkle(wg, oracle, h):
    from starkpack_winterfell_b200.merkle import MerkleTree

    for leaves in (fx.LEAVES4, fx.LEAVES8):  # crypto/src/merkle/tests.rs:12-99
        lv = np.array(leaves, dtype=np.uint8)
        t = MerkleTree.new(h, lv)
        want = oracle.merkle_nodes(h, lv)
        assert np.array_equal(t.nodes, want)
        assert t.root() == bytes(want[1]) and t.depth() == len(leaves).bit_length() - 1
        # root == nested merge, as the reference's test states it
        mg = lambda a, b: oracle.hash_bytes(h, bytes(a) + bytes(b))
        lvl = [bytes(x) for x in lv]
        while len(lvl) > 1:
            lvl = [mg(lvl[i], lvl[i + 1]) for i in range(0, len(lvl), 2)]
        assert t.root() == lvl[0]
    rng = np.random.default_rng(7)
    for n in [2, 4, 16, 128, 1024, 2048, 4096, 1 << 15, 1 << 18]:
        lv = rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
        t = MerkleTree.new(h, lv)
        assert np.array_equal(t.nodes, oracle.merkle_nodes(h, lv)), n
        assert not t.nodes[0].any()


def _commit(wg, field, h, traces, blowup):
    cms = [wg.matrix.ColMatrix(field, list(t)) for t in traces]
    dom = wg.matrix.StarkDomain(field, cms[0].num_rows(), blowup)
    return wg.prover.build_trace_commitment(h, cms, dom)


@pytest.mark.parametrize("field,h,shapes,n,blowup", [
    (F128, B3, [2], 8, 8),                  # prover/src/trace/tests.rs:42-128 fib trace
    (F64, B3, [64], 256, 8),
    (F64, B3, [100], 2048, 8),              # cfg2 shape, small n
    (F64, B3, [10], 8192, 8),               # two-pass NTT
    (F128, SHA3, [4], 4096, 8),             # cfg3 shape
    (F128, B3, [2] * 8, 1024, 8),           # cfg4: StarkPack aggregation of 8 traces
    (F128, B3, [2] * 8, 8192, 8),
    (F64, SHA3, [3, 5, 9], 512, 4),         # ragged aggregation
    (F64, B3, [256], 1024, 8),              # cfg5 shape: 2048-byte rows = 2 chunks + parent
    (F64, B3, [130, 130], 256, 2),
])
def test_commit_trace(wg, oracle, field, h, shapes, n, blowup):
    if field == F128 and shapes == [2] and n == 8:
        traces = [np.array([[(v, 0) for v in col] for col in fx.FIB_COLS], dtype=np.uint64)]
    elif field == F128 and len(shapes) == 8:
        traces = [fx.fib_trace_f128(n, (t + 1, t + 1)) for t in range(8)]  # examples/src/do_work/mod.rs:56-60
    else:
        traces = [rand_cols(field, 500 + 13 * i + c, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, h, traces, blowup)
    ldes, tree, polys = _commit(wg, field, h, traces, blowup)
    for t in range(len(traces)):
        for c in range(shapes[t]):
            assert np.array_equal(polys[t].columns[c], want["polys"][t][c])
        assert np.array_equal(ldes[t].data, want["lde"][t])
    assert np.array_equal(tree.leaves(), want["leaves"])
    assert np.array_equal(tree.nodes, want["nodes"])
    assert tree.root() == bytes(want["root"])
    tc = wg.prover.TraceCommitment(ldes, tree, [blowup] * len(traces))
    assert tc.root() == tree.root()


def test_lde_interpolates_back(wg):
    """prover/src/trace/tests.rs:88-104: LDE columns interpolated over the shifted domain give back the polys
    (checked here through the GPU path itself: un-shift by powers of offset^-1)."""
    n, blowup = 64, 8
    tr = fx.synth_f128(9, 2, n)
    polys = wg.matrix.ColMatrix(F128, list(tr)).interpolate_columns()
    lde = wg.matrix.RowMatrix.evaluate_polys(polys, blowup)
    q = fx.P128
    for c in range(2):
        col = np.ascontiguousarray(lde.data[:, c, :]).copy()
        wg.fft.interpolate_poly(F128, col)  # coefficients of P(offset * x): c_m * offset^m, degree < n
        inv = pow(3, -1, q)
        for m in range(n * blowup):
            v = (int(col[m, 0]) | (int(col[m, 1]) << 64)) * pow(inv, m, q) % q
            want = (int(polys.columns[c][m, 0]) | (int(polys.columns[c][m, 1]) << 64)) if m < n else 0
            assert v == want


def test_error_codes(wg):
    from starkpack_winterfell_b200._lib import ERR_NOT_POW2, ERR_TOO_LARGE, ERR_ZERO_OFFSET

    with pytest.raises(wg.WinterGpuError) as e:  # fft/mod.rs:279-282
        wg.fft.interpolate_poly(F64, np.zeros(12, dtype=np.uint64))
    assert e.value.code == ERR_NOT_POW2
    with pytest.raises(wg.WinterGpuError) as e:  # fft/mod.rs:197-200
        wg.fft.evaluate_poly_with_offset(F64, np.zeros(8, dtype=np.uint64), None, np.zeros(1, dtype=np.uint64), 4)
    assert e.value.code == ERR_ZERO_OFFSET
    with pytest.raises(wg.WinterGpuError) as e:  # blowup not a power of two, fft/mod.rs:185-188
        wg.fft.evaluate_poly_with_offset(F64, np.zeros(8, dtype=np.uint64), None, np.ones(1, dtype=np.uint64), 3)
    assert e.value.code == ERR_NOT_POW2
    with pytest.raises(wg.WinterGpuError) as e:
        wg.fft.get_twiddles(F64, 1 << 33)
    assert e.value.code == ERR_TOO_LARGE
    with pytest.raises(ValueError):  # twiddle length contract, fft/mod.rs:283-289
        wg.fft.interpolate_poly(F64, np.zeros(8, dtype=np.uint64), np.zeros(3, dtype=np.uint64))
