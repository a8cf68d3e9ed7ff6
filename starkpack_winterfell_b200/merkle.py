"""Mirror of winter_crypto::MerkleTree construction (crypto/src/merkle/mod.rs:87-182) over the C ABI."""
import ctypes as C

import numpy as np

from ._lib import BLAKE3_256, SHA3_256, WinterGpuError, check, lib  # noqa: F401


class MerkleTreeError(ValueError):
    """crypto/src/errors.rs:13-26"""


class MerkleTree:
    def __init__(self, hash_id, nodes, leaves):
        self.hash_id, self.nodes, self._leaves = hash_id, nodes, leaves

    @classmethod
    def new(cls, hash_id, leaves):
        """MerkleTree::new (crypto/src/merkle/mod.rs:117-136)."""
        leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, 32)
        n = leaves.shape[0]
        if n < 2:
            raise MerkleTreeError(f"TooFewLeaves(2, {n})")
        if n & (n - 1):
            raise MerkleTreeError(f"NumberOfLeavesNotPowerOfTwo({n})")
        nodes = np.empty((n, 32), dtype=np.uint8)
        check(lib().wg_merkle_new(hash_id, C.c_void_p(leaves.ctypes.data), n, C.c_void_p(nodes.ctypes.data)))
        return cls(hash_id, nodes, leaves)

    @classmethod
    def from_raw_parts(cls, hash_id, nodes, leaves):
        """MerkleTree::from_raw_parts (crypto/src/merkle/mod.rs:149-161)."""
        nodes = np.ascontiguousarray(nodes, dtype=np.uint8).reshape(-1, 32)
        leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, 32)
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
