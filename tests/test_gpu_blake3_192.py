"""Blake3_192 (crypto/src/hash/blake/mod.rs:68-114; hash id 2 of the C ABI): BLAKE3 truncated to 24 bytes, merge = hash of
the 48 bytes of two truncated digests. Leaves, nodes and roots are 24 bytes. Checked against the oracle (itself pinned
against the Python `blake3` module in tests/test_oracle_hash.py) and directly against `blake3`."""
import blake3
import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu
H192 = 2


@pytest.fixture(scope="module")
def wg():
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    assert w._lib.lib().wg_digest_bytes(H192) == 24 and w._lib.lib().wg_digest_bytes(0) == 32 and w._lib.lib().wg_digest_bytes(7) == 0
    return w


@pytest.mark.parametrize("field,cols", [(0, 1), (0, 8), (0, 17), (0, 100), (0, 128), (0, 129), (0, 300), (1, 1), (1, 4), (1, 64), (1, 65)])
def test_hash_rows(wg, oracle, field, cols):
    rows, rw = 77, (cols + 7) // 8 * 8
    m = fx.synth_f64(cols, rw, rows).T.copy() if field == 0 else np.ascontiguousarray(fx.synth_f128(cols, rw, rows).transpose(1, 0, 2))
    got = wg.matrix.RowMatrix(field, m, rw, cols).hash_rows(H192)
    assert got.shape == (rows, 24)
    assert np.array_equal(got, oracle.hash_rows(field, H192, m, rows, rw, cols))
    if field == 1:  # canonical field: the hashed bytes are the raw row memory
        assert bytes(got[5]) == blake3.blake3(m[5, :cols].tobytes()).digest()[:24]


@pytest.mark.parametrize("n", [2, 4, 8, 1024, 4096, 1 << 15])
def test_merkle_tree(wg, oracle, n):
    rng = np.random.default_rng(n)
    leaves = rng.integers(0, 256, (n, 24), dtype=np.uint8)
    tree = wg.merkle.MerkleTree.new(H192, leaves)
    assert tree.nodes.shape == (n, 24) and not tree.nodes[0].any()
    assert np.array_equal(tree.nodes, oracle.merkle_nodes(H192, leaves))
    h = lambda a, b: blake3.blake3(bytes(a) + bytes(b)).digest()[:24]  # noqa: E731
    assert bytes(tree.nodes[n // 2]) == h(leaves[0], leaves[1])
    if n == 4:
        assert tree.root() == h(h(leaves[0], leaves[1]), h(leaves[2], leaves[3]))


@pytest.mark.parametrize("field,shapes,n,blowup", [(0, [21], 4096, 8), (0, [100], 1 << 13, 8), (1, [2, 2, 2], 512, 8), (0, [300], 256, 2)])
def test_commit_trace(wg, oracle, field, shapes, n, blowup):
    traces = [fx.synth_f64(31 + i, c, n) if field == 0 else fx.synth_f128(31 + i, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, H192, traces, blowup, want=("leaves", "nodes"))
    cms = [wg.matrix.ColMatrix(field, list(t)) for t in traces]
    ldes, tree, polys = wg.prover.build_trace_commitment(H192, cms, wg.matrix.StarkDomain(field, n, blowup))
    assert tree.root() == bytes(want["root"]) and len(tree.root()) == 24
    assert np.array_equal(tree.leaves(), want["leaves"]) and np.array_equal(tree.nodes, want["nodes"])
    tree2 = ldes[0].commit_to_comb_rows(H192, ldes)  # over the already-built LDE matrices
    assert tree2.root() == tree.root()
    rc = wg.prover.ResidentCommitment.from_traces(H192, cms, wg.matrix.StarkDomain(field, n, blowup))
    assert rc.root() == tree.root()
    pos = [0, 5, n * blowup - 1]
    proof = rc.prove_batch(pos)
    host = tree.prove_batch(pos)
    assert [bytes(x) for x in proof.leaves] == [bytes(x) for x in host.leaves]
    assert [[bytes(y) for y in x] for x in proof.nodes] == [[bytes(y) for y in x] for x in host.nodes]
    rc.close()


def test_sharded_entry_rejects_blake3_192(wg):
    import torch

    from starkpack_winterfell_b200 import _lib, sharded

    d = torch.zeros((4, 64, 1), dtype=torch.int64, device="cuda:0")
    with pytest.raises(_lib.WinterGpuError):
        sharded.commit_emulated(0, H192, 64, 8, 4, 2, d)
