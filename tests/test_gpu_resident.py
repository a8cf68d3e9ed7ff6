"""GPU parity of the resident commitment (LDE + tree kept in HBM, queried by position) and of the constraint commitment
(Prover::build_constraint_commitment), against the CPU oracle. Mirrors prover/src/trace/commitment.rs:87-190 (query),
crypto/src/merkle/mod.rs:194-284 (prove / prove_batch) and prover/src/lib.rs:680-715."""
import numpy as np
import pytest

from tests import fixtures as fx
from tests.proof_check import batch_root, path_root

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


def rand_cols(field, seed, cols, n):
    return fx.synth_f64(seed, cols, n) if field == F64 else fx.synth_f128(seed, cols, n)


def col_matrices(wg, field, traces):
    return [wg.matrix.ColMatrix(field, [np.ascontiguousarray(c) for c in t]) for t in traces]


@pytest.mark.parametrize("field,h,shapes,n,blowup", [
    (F64, B3, [13], 4096, 8),
    (F64, SHA3, [3, 5], 1 << 13, 4),
    (F128, B3, [2, 2, 2, 2], 1024, 8),
    (F128, SHA3, [4], 1 << 13, 8),
    (F64, B3, [100], 1 << 14, 8),
    (F64, 2, [13], 4096, 8),          # Blake3_192: 24-byte digests through leaves, nodes, proofs
    (F128, 2, [2, 2], 1024, 8),
])
def test_resident_commitment_query(wg, oracle, field, h, shapes, n, blowup):
    traces = [rand_cols(field, 700 + 7 * i + c, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, h, traces, blowup)
    dom = wg.matrix.StarkDomain(field, n, blowup)
    rc = wg.prover.ResidentCommitment.from_traces(h, col_matrices(wg, field, traces), dom)
    L = n * blowup
    assert rc.root() == bytes(want["root"]) and rc.lde_rows == L and rc.cols_per_trace == shapes
    for t in range(len(shapes)):
        for c in range(shapes[t]):
            assert np.array_equal(rc.polys[t].columns[c], want["polys"][t][c])
    rng = np.random.default_rng(n + len(shapes))
    merge = lambda a, b: oracle.hash_bytes(h, a + b)  # noqa: E731
    host_tree = wg.merkle.MerkleTree.from_raw_parts(h, want["nodes"], want["leaves"])
    for k in (1, 32, 255):
        pos = [int(x) for x in rng.choice(L, size=k, replace=False)]
        q = rc.query(pos)
        for t in range(len(shapes)):
            assert np.array_equal(q.traces_states[t], want["lde"][t][pos][:, :shapes[t]]), (t, k)
        assert q.comb_states.shape[1] == sum(shapes)
        assert q.trace_proof == host_tree.prove_batch(pos)
        assert batch_root(merge, q.trace_proof, pos) == rc.root()
        # the leaves of the proof are the hashes of the concatenated rows (row_matrix.rs:217-231)
        got_leaf = oracle.hash_comb_rows(field, h, [np.ascontiguousarray(s) for s in q.traces_states], shapes)
        assert [bytes(x) for x in got_leaf] == [bytes(x) for x in q.trace_proof.leaves]
    p = rc.prove(pos[0])
    assert [bytes(x) for x in p] == [bytes(x) for x in host_tree.prove(pos[0])]
    assert path_root(merge, pos[0], p) == rc.root()
    with pytest.raises(wg.merkle.MerkleTreeError):
        rc.prove_batch([0, 0])
    with pytest.raises(wg.merkle.MerkleTreeError):
        rc.prove_batch([L])
    with pytest.raises(wg.WinterGpuError):
        rc.rows(0, [L])
    mats, tree = rc.download()
    for t in range(len(shapes)):
        assert np.array_equal(mats[t].data, want["lde"][t])
    assert np.array_equal(tree.nodes, want["nodes"]) and np.array_equal(tree.leaves(), want["leaves"])
    rc.close()


@pytest.mark.parametrize("field,h,cols,n,blowup", [(F64, B3, 5, 2048, 8), (F128, SHA3, 3, 1 << 13, 4), (F64, B3, 8, 1 << 14, 8)])
def test_constraint_commitment(wg, oracle, field, h, cols, n, blowup):
    """build_constraint_commitment(polys) == commitment of the trace those polys interpolate."""
    polys = rand_cols(field, 900 + cols, cols, n)
    one = np.array([0xFFFFFFFF], dtype=np.uint64) if field == F64 else np.array([1, 0], dtype=np.uint64)
    trace = np.stack([oracle.evaluate_poly_with_offset(field, polys[c], one, 1) for c in range(cols)])
    want = oracle.commit_trace(field, h, [trace], blowup)
    assert np.array_equal(np.stack(want["polys"][0]), polys)
    dom = wg.matrix.StarkDomain(field, n, blowup)
    cm = wg.matrix.ColMatrix(field, [np.ascontiguousarray(c) for c in polys])
    lde, tree = wg.prover.build_constraint_commitment(h, cm, dom)
    assert np.array_equal(lde.data, want["lde"][0])
    assert np.array_equal(tree.nodes, want["nodes"]) and tree.root() == bytes(want["root"])
    rc = wg.prover.ResidentCommitment.from_polys(h, cm, dom)
    assert rc.root() == bytes(want["root"])
    pos = [1, 2, n * blowup - 1, 77]
    assert np.array_equal(rc.rows(0, pos), want["lde"][0][pos][:, :cols])
    rc.close()
