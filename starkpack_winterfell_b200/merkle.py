"""Mirror of winter_crypto::MerkleTree construction (crypto/src/merkle/mod.rs:87-182) over the C ABI."""
import ctypes as C

import numpy as np

from ._lib import BLAKE3_256, DIGEST_BYTES, SHA3_256, WinterGpuError, check, lib  # noqa: F401


class MerkleTreeError(ValueError):
    """crypto/src/errors.rs:13-26"""


MAX_PATHS = 255  # crypto/src/merkle/proofs.rs:18


class BatchMerkleProof:
    """crypto/src/merkle/proofs.rs:31-38: the queried leaves (in query order), per normalized index the sibling digests
    needed bottom-up, and the tree depth."""

    def __init__(self, leaves, nodes, depth):
        self.leaves, self.nodes, self.depth = leaves, nodes, depth

    def __eq__(self, o):
        return (self.depth == o.depth and [bytes(x) for x in self.leaves] == [bytes(x) for x in o.leaves]
                and [[bytes(x) for x in v] for v in self.nodes] == [[bytes(x) for x in v] for v in o.nodes])


def map_indexes(indexes, depth):
    """crypto/src/merkle/mod.rs:376-395"""
    num_leaves = 1 << depth
    m = {}
    for i, index in enumerate(indexes):
        m[index] = i
        if index >= num_leaves:
            raise MerkleTreeError(f"LeafIndexOutOfBounds({num_leaves}, {index})")
    if len(m) != len(indexes):
        raise MerkleTreeError("DuplicateLeafIndex")
    return m


def normalize_indexes(indexes):
    """crypto/src/merkle/mod.rs:397-403"""
    return sorted({i - (i & 1) for i in indexes})


class _DigestPlan:
    """First pass of prove / prove_batch over a tree whose digests are not in host memory: records which leaves and
    nodes the reference algorithm reads, so that they can be fetched in one gather."""

    def __init__(self):
        self.leaf_idx, self.node_idx = [], []

    def leaf(self, i):
        self.leaf_idx.append(i)
        return None

    def node(self, i):
        self.node_idx.append(i)
        return None


class _DigestTable:
    def __init__(self, leaf_idx, leaf_vals, node_idx, node_vals):
        self.l = {i: leaf_vals[k] for k, i in enumerate(leaf_idx)}
        self.n = {i: node_vals[k] for k, i in enumerate(node_idx)}

    def leaf(self, i):
        return self.l[i]

    def node(self, i):
        return self.n[i]


def _prove(src, n_leaves, index):
    """MerkleTree::prove (crypto/src/merkle/mod.rs:194-211)"""
    if index >= n_leaves:
        raise MerkleTreeError(f"LeafIndexOutOfBounds({n_leaves}, {index})")
    proof = [src.leaf(index), src.leaf(index ^ 1)]
    index = (index + n_leaves) >> 1
    while index > 1:
        proof.append(src.node(index ^ 1))
        index >>= 1
    return proof


def _prove_batch(src, n_leaves, depth, indexes):
    """MerkleTree::prove_batch (crypto/src/merkle/mod.rs:222-284)"""
    indexes = [int(i) for i in indexes]
    if not indexes:
        raise MerkleTreeError("TooFewLeafIndexes")
    if len(indexes) > MAX_PATHS:
        raise MerkleTreeError(f"TooManyLeafIndexes({MAX_PATHS}, {len(indexes)})")
    index_map = map_indexes(indexes, depth)
    norm = normalize_indexes(indexes)
    leaves = [None] * len(index_map)
    nodes = []
    next_indexes = []
    for index in norm:
        missing = []
        for i in (index, index + 1):
            v = src.leaf(i)
            if i in index_map:
                leaves[index_map[i]] = v
            else:
                missing.append(v)
        nodes.append(missing)
        next_indexes.append((index + n_leaves) >> 1)
    for _ in range(1, depth):
        cur, next_indexes = next_indexes, []
        i = 0
        while i < len(cur):
            sibling = cur[i] ^ 1
            if i + 1 < len(cur) and cur[i + 1] == sibling:
                i += 1
            else:
                nodes[i].append(src.node(sibling))
            next_indexes.append(sibling >> 1)
            i += 1
    return BatchMerkleProof(leaves, nodes, depth)


class MerkleTree:
    def __init__(self, hash_id, nodes, leaves):
        self.hash_id, self.nodes, self._leaves = hash_id, nodes, leaves

    # digest source over the host arrays
    def leaf(self, i):
        return self._leaves[i]

    def node(self, i):
        return self.nodes[i]

    def prove(self, index):
        """MerkleTree::prove (crypto/src/merkle/mod.rs:194-211): [leaf, sibling leaf, siblings bottom-up]."""
        return _prove(self, self._leaves.shape[0], int(index))

    def prove_batch(self, indexes):
        """MerkleTree::prove_batch (crypto/src/merkle/mod.rs:222-284)."""
        return _prove_batch(self, self._leaves.shape[0], self.depth(), indexes)

    @classmethod
    def new(cls, hash_id, leaves):
        """MerkleTree::new (crypto/src/merkle/mod.rs:117-136)."""
        leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, DIGEST_BYTES[hash_id])
        n = leaves.shape[0]
        if n < 2:
            raise MerkleTreeError(f"TooFewLeaves(2, {n})")
        if n & (n - 1):
            raise MerkleTreeError(f"NumberOfLeavesNotPowerOfTwo({n})")
        nodes = np.empty((n, DIGEST_BYTES[hash_id]), dtype=np.uint8)
        check(lib().wg_merkle_new(hash_id, C.c_void_p(leaves.ctypes.data), n, C.c_void_p(nodes.ctypes.data)))
        return cls(hash_id, nodes, leaves)

    @classmethod
    def from_raw_parts(cls, hash_id, nodes, leaves):
        """MerkleTree::from_raw_parts (crypto/src/merkle/mod.rs:149-161)."""
        nodes = np.ascontiguousarray(nodes, dtype=np.uint8).reshape(-1, DIGEST_BYTES[hash_id])
        leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, DIGEST_BYTES[hash_id])
        if leaves.shape[0] < 2 or leaves.shape[0] & (leaves.shape[0] - 1):
            raise MerkleTreeError("number of leaves must be a power of two >= 2")
        if nodes.shape[0] != leaves.shape[0]:
            raise MerkleTreeError("number of leaves must be equal to number of nodes")
        return cls(hash_id, nodes, leaves)

    def root(self):
        return bytes(self.nodes[1])

    def leaves(self):
        return self._leaves

    def depth(self):
        return self._leaves.shape[0].bit_length() - 1
