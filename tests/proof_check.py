"""Test-side verifier for Merkle proofs: restates MerkleTree::verify (crypto/src/merkle/mod.rs:293-312) and
BatchMerkleProof::get_root (crypto/src/merkle/proofs.rs:140-274) with the oracle's hash as `merge`."""
from starkpack_winterfell_b200.merkle import map_indexes, normalize_indexes


def path_root(merge, index, proof):
    r = index & 1
    v = merge(bytes(proof[r]), bytes(proof[1 - r]))
    index = (index + (1 << (len(proof) - 1))) >> 1
    for p in proof[2:]:
        v = merge(v, bytes(p)) if index & 1 == 0 else merge(bytes(p), v)
        index >>= 1
    return v


def batch_root(merge, proof, indexes):
    indexes = [int(i) for i in indexes]
    index_map = map_indexes(indexes, proof.depth)
    norm = normalize_indexes(indexes)
    assert len(norm) == len(proof.nodes)
    offset = 1 << proof.depth
    v, nxt, ptr = {}, [], []
    for i, index in enumerate(norm):
        if index in index_map:
            a = bytes(proof.leaves[index_map[index]])
            if index + 1 in index_map:
                b = bytes(proof.leaves[index_map[index + 1]])
                ptr.append(0)
            else:
                b = bytes(proof.nodes[i][0])
                ptr.append(1)
        else:
            a = bytes(proof.nodes[i][0])
            b = bytes(proof.leaves[index_map[index + 1]])
            ptr.append(1)
        parent = (offset + index) >> 1
        v[parent] = merge(a, b)
        nxt.append(parent)
    for _ in range(1, proof.depth):
        cur, nxt = nxt, []
        i = 0
        while i < len(cur):
            node_index = cur[i]
            sib_index = node_index ^ 1
            if i + 1 < len(cur) and cur[i + 1] == sib_index:
                sib = v[sib_index]
                i += 1
            else:
                sib = bytes(proof.nodes[i][ptr[i]])
                ptr[i] += 1
            node = v[node_index]
            parent = merge(sib, node) if node_index & 1 else merge(node, sib)
            v[node_index >> 1] = parent
            nxt.append(node_index >> 1)
            i += 1
    return v[1]
