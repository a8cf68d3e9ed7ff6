"""Host logic of MerkleTree::prove / prove_batch (starkpack_winterfell_b200/merkle.py) against the reference's own
expectations (crypto/src/merkle/tests.rs:94-233), on trees built by the CPU oracle. No GPU needed."""
import numpy as np
import pytest

import oracle
from starkpack_winterfell_b200.merkle import MerkleTree, MerkleTreeError
from tests import fixtures as fx
from tests.proof_check import batch_root, path_root

B3, SHA3 = 0, 1


def merge(h):
    return lambda a, b: oracle.hash_bytes(h, a + b)


def tree_of(h, leaves):
    leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, 32)
    return MerkleTree.from_raw_parts(h, oracle.merkle_nodes(h, leaves), leaves)


def b(x):
    return [bytes(v) for v in x]


def test_prove_matches_reference_expectations():
    h2 = merge(B3)
    l4 = b(np.array(fx.LEAVES4, dtype=np.uint8).reshape(-1, 32))
    t = tree_of(B3, fx.LEAVES4)
    assert b(t.prove(1)) == [l4[1], l4[0], h2(l4[2], l4[3])]
    assert b(t.prove(2)) == [l4[2], l4[3], h2(l4[0], l4[1])]
    l8 = b(np.array(fx.LEAVES8, dtype=np.uint8).reshape(-1, 32))
    t = tree_of(B3, fx.LEAVES8)
    assert b(t.prove(1)) == [l8[1], l8[0], h2(l8[2], l8[3]), h2(h2(l8[4], l8[5]), h2(l8[6], l8[7]))]
    assert b(t.prove(6)) == [l8[6], l8[7], h2(l8[4], l8[5]), h2(h2(l8[0], l8[1]), h2(l8[2], l8[3]))]
    for i in (1, 6):
        assert path_root(h2, i, t.prove(i)) == t.root()
    with pytest.raises(MerkleTreeError):
        t.prove(8)


def test_prove_batch_matches_reference_expectations():
    h2 = merge(B3)
    l8 = b(np.array(fx.LEAVES8, dtype=np.uint8).reshape(-1, 32))
    t = tree_of(B3, fx.LEAVES8)
    p = t.prove_batch([1])
    assert b(p.leaves) == [l8[1]] and p.depth == 3
    assert [b(v) for v in p.nodes] == [[l8[0], h2(l8[2], l8[3]), h2(h2(l8[4], l8[5]), h2(l8[6], l8[7]))]]
    p = t.prove_batch([1, 2])
    assert b(p.leaves) == [l8[1], l8[2]]
    assert [b(v) for v in p.nodes] == [[l8[0], h2(h2(l8[4], l8[5]), h2(l8[6], l8[7]))], [l8[3]]]
    p = t.prove_batch([1, 6])
    assert b(p.leaves) == [l8[1], l8[6]]
    assert [b(v) for v in p.nodes] == [[l8[0], h2(l8[2], l8[3])], [l8[7], h2(l8[4], l8[5])]]
    p = t.prove_batch(list(range(8)))
    assert b(p.leaves) == l8 and [b(v) for v in p.nodes] == [[], [], [], []]
    # verify_batch shapes of the reference (tests.rs:210-233)
    for idx in ([1], [1, 2], [1, 6], [1, 3, 6], [0, 1, 2, 3, 4, 5, 6, 7], [6, 1], [7, 2, 5]):
        assert batch_root(h2, t.prove_batch(idx), idx) == t.root(), idx


def test_prove_batch_errors():
    t = tree_of(B3, fx.LEAVES8)
    with pytest.raises(MerkleTreeError, match="TooFewLeafIndexes"):
        t.prove_batch([])
    with pytest.raises(MerkleTreeError, match="DuplicateLeafIndex"):
        t.prove_batch([1, 1])
    with pytest.raises(MerkleTreeError, match="LeafIndexOutOfBounds"):
        t.prove_batch([8])
    big = tree_of(SHA3, np.arange(512 * 32, dtype=np.uint32).astype(np.uint8).reshape(512, 32))
    with pytest.raises(MerkleTreeError, match="TooManyLeafIndexes"):
        big.prove_batch(list(range(256)))


@pytest.mark.parametrize("h", [B3, SHA3])
def test_random_batches_resolve_to_root(h):
    rng = np.random.default_rng(5)
    leaves = rng.integers(0, 256, (1024, 32), dtype=np.uint8)
    t = tree_of(h, leaves)
    for k in (1, 2, 7, 33, 255):
        idx = [int(x) for x in rng.choice(1024, size=k, replace=False)]
        assert batch_root(merge(h), t.prove_batch(idx), idx) == t.root()
        assert path_root(merge(h), idx[0], t.prove(idx[0])) == t.root()
