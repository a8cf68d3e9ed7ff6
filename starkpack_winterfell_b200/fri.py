"""Mirror of the FRI prover's commit phase (fri/src/prover/mod.rs:100-330, fri/src/folding/mod.rs, fri/src/options.rs)
over the C ABI — SURVEY 8(f) rank 3. Base field only (E = B, FieldExtension::None).

What runs on the GPU: per layer the transpose + row hashing + Merkle tree (wg_fri_commit_layer[_dev]) and the
degree-respecting projection (wg_fri_apply_drp[_dev]); the remainder interpolation (wg_interpolate_poly_with_offset) and
its digest (wg_hash_rows). What stays with the caller: the Fiat-Shamir channel — any object with
`commit_fri_layer(root: bytes)` and `draw_fri_alpha() -> element` (internal representation: f64 Montgomery word as
uint64[1], f128 canonical as uint64[2]), i.e. the reference's ProverChannel trait (fri/src/prover/channel.rs:22-44).
"""
import ctypes as C

import numpy as np

from . import fft
from ._lib import DIGEST_BYTES, ELEM_BYTES, F64, check, lib
from .matrix import GENERATOR_INTERNAL
from .merkle import MerkleTree, _DigestPlan, _DigestTable, _prove_batch


class FriOptions:
    """fri/src/options.rs:14-94"""

    def __init__(self, blowup_factor, folding_factor, remainder_max_degree):
        assert blowup_factor & (blowup_factor - 1) == 0, f"blowup factor must be a power of two, but was {blowup_factor}"
        assert folding_factor in (2, 4, 8, 16), f"folding factor {folding_factor} is not supported"
        self.blowup_factor, self.folding_factor, self.remainder_max_degree = blowup_factor, folding_factor, remainder_max_degree

    def domain_offset(self, field):
        return GENERATOR_INTERNAL[field]  # options.rs:57-59: B::GENERATOR

    def num_fri_layers(self, domain_size):
        """options.rs:85-93"""
        result = 0
        max_remainder_size = (self.remainder_max_degree + 1) * self.blowup_factor
        while domain_size > max_remainder_size:
            domain_size //= self.folding_factor
            result += 1
        return result


def fold_positions(positions, source_domain_size, folding_factor):
    """fri/src/folding/mod.rs:160-176"""
    target = source_domain_size // folding_factor
    result = []
    for p in positions:
        p = int(p) % target
        if p not in result:
            result.append(p)
    return result


def _w(field):
    return ELEM_BYTES[field] // 8


def commit_layer(field, hash_id, evaluations, folding_factor):
    """transpose_slice + hash_values + MerkleTree::new (fri/src/prover/mod.rs:191-199), host buffers.
    Returns (transposed [rows, N(,2)], MerkleTree)."""
    e = np.ascontiguousarray(evaluations, dtype=np.uint64)
    n = e.shape[0]
    rows = n // folding_factor
    tr = np.empty_like(e)
    leaves = np.empty((max(rows, 1), DIGEST_BYTES[hash_id]), dtype=np.uint8)
    nodes = np.empty((max(rows, 1), DIGEST_BYTES[hash_id]), dtype=np.uint8)
    check(lib().wg_fri_commit_layer(field, hash_id, C.c_void_p(e.ctypes.data), n, folding_factor, C.c_void_p(tr.ctypes.data),
                                    C.c_void_p(leaves.ctypes.data), C.c_void_p(nodes.ctypes.data)))
    shape = (rows, folding_factor) if field == F64 else (rows, folding_factor, 2)
    return tr.reshape(shape), MerkleTree.from_raw_parts(hash_id, nodes, leaves)


def apply_drp(field, transposed, domain_offset, alpha):
    """folding::apply_drp (fri/src/folding/mod.rs:115-146), host buffers. transposed [rows, N(,2)] -> [rows(,2)]."""
    t = np.ascontiguousarray(transposed, dtype=np.uint64)
    rows, n = t.shape[0], t.shape[1]
    out = np.empty((rows,) if field == F64 else (rows, 2), dtype=np.uint64)
    off = np.ascontiguousarray(domain_offset, dtype=np.uint64)
    al = np.ascontiguousarray(alpha, dtype=np.uint64)
    check(lib().wg_fri_apply_drp(field, C.c_void_p(t.ctypes.data), rows, n, C.c_void_p(off.ctypes.data), C.c_void_p(al.ctypes.data),
                                 C.c_void_p(out.ctypes.data)))
    return out


class FriLayer:
    """fri/src/prover/mod.rs:110-114, buffers resident in HBM (torch tensors)."""

    def __init__(self, hash_id, evaluations, leaves, nodes, folding):
        self.hash_id, self.evaluations, self.leaves, self.nodes, self.folding = hash_id, evaluations, leaves, nodes, folding

    def root(self):
        return bytes(self.nodes[1].cpu().numpy())

    def tree(self):
        """full host copy as a MerkleTree (tests; the proof path below only gathers what it reads)."""
        return MerkleTree.from_raw_parts(self.hash_id, self.nodes.cpu().numpy(), self.leaves.cpu().numpy())


class FriProofLayer:
    """fri/src/proof.rs FriProofLayer before serialisation: the queried [E; N] values and the batch Merkle proof."""

    def __init__(self, values, proof):
        self.values, self.proof = values, proof


class FriProver:
    """fri/src/prover/mod.rs:100-330 with every layer kept on the device."""

    def __init__(self, field, hash_id, options, device="cuda:0"):
        import torch

        self.field, self.hash_id, self.options = field, hash_id, options
        self.device = torch.device(device)
        self.layers = []
        self.remainder_poly = None
        self.remainder_commitment = None

    def folding_factor(self):
        return self.options.folding_factor

    def domain_offset(self):
        return self.options.domain_offset(self.field)

    def num_layers(self):
        return len(self.layers)

    def reset(self):
        self.layers = []
        self.remainder_poly = None

    # ---- commit phase -------------------------------------------------------------------------------------------
    def build_layers(self, channel, evaluations):
        """prover/mod.rs:171-190. `evaluations`: numpy array or torch device tensor [n(,2)] (uint64 / int64 words)."""
        import torch

        assert not self.layers, "a prior proof generation request has not been completed yet"
        if isinstance(evaluations, torch.Tensor):
            ev = evaluations.to(self.device).contiguous().view(torch.int64)
        else:
            ev = torch.from_numpy(np.ascontiguousarray(evaluations, dtype=np.uint64).view(np.int64)).to(self.device)
        ev = ev.reshape(-1, _w(self.field))
        for _ in range(self.options.num_fri_layers(ev.shape[0])):
            ev = self._build_layer(channel, ev)
        self._set_remainder(channel, ev)

    def _build_layer(self, channel, ev):
        """prover/mod.rs:192-211"""
        import torch

        n, N, w = ev.shape[0], self.folding_factor(), _w(self.field)
        rows = n // N
        s = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        tr = torch.empty((rows, N, w), dtype=torch.int64, device=self.device)
        leaves = torch.empty((max(rows, 1), DIGEST_BYTES[self.hash_id]), dtype=torch.uint8, device=self.device)
        nodes = torch.empty((max(rows, 1), DIGEST_BYTES[self.hash_id]), dtype=torch.uint8, device=self.device)
        check(lib().wg_fri_commit_layer_dev(self.field, self.hash_id, C.c_void_p(ev.data_ptr()), n, N, C.c_void_p(tr.data_ptr()),
                                            C.c_void_p(leaves.data_ptr()), C.c_void_p(nodes.data_ptr()), s))
        layer = FriLayer(self.hash_id, tr, leaves, nodes, N)
        channel.commit_fri_layer(layer.root())  # 32 bytes D2H: the only transfer of the layer
        alpha = np.ascontiguousarray(channel.draw_fri_alpha(), dtype=np.uint64)
        off = np.ascontiguousarray(self.domain_offset(), dtype=np.uint64)
        nxt = torch.empty((rows, w), dtype=torch.int64, device=self.device)
        check(lib().wg_fri_apply_drp_dev(self.field, C.c_void_p(tr.data_ptr()), rows, N, C.c_void_p(off.ctypes.data),
                                         C.c_void_p(alpha.ctypes.data), C.c_void_p(nxt.data_ptr()), s))
        self.layers.append(layer)
        return nxt

    def _set_remainder(self, channel, ev):
        """prover/mod.rs:214-223"""
        e = np.ascontiguousarray(ev.cpu().numpy().view(np.uint64))
        e = e.reshape(-1) if self.field == F64 else e.reshape(-1, 2)
        n = e.shape[0]
        if n > 1:
            fft.interpolate_poly_with_offset(self.field, e, fft.get_inv_twiddles(self.field, n), self.domain_offset())
        size = n // self.options.blowup_factor
        rem = np.ascontiguousarray(e[:size])
        digest = np.empty((1, DIGEST_BYTES[self.hash_id]), dtype=np.uint8)
        check(lib().wg_hash_rows(self.field, self.hash_id, C.c_void_p(rem.ctypes.data), 1, size, size, C.c_void_p(digest.ctypes.data)))
        self.remainder_commitment = bytes(digest[0])
        channel.commit_fri_layer(self.remainder_commitment)
        self.remainder_poly = rem

    # ---- query phase --------------------------------------------------------------------------------------------
    def build_proof(self, positions):
        """prover/mod.rs:236-279: per layer the folded positions' [E; N] rows and the batch Merkle proof; then the
        remainder polynomial. Digests and rows are gathered on the device, only the queried ones cross PCIe."""
        import torch

        assert self.remainder_poly is not None, "FRI layers have not been built yet"
        out = []
        positions = [int(p) for p in positions]
        if self.layers:
            domain_size = self.layers[0].evaluations.shape[0] * self.folding_factor()
            for layer in self.layers:
                positions = fold_positions(positions, domain_size, self.folding_factor())
                n_leaves = layer.leaves.shape[0]
                depth = n_leaves.bit_length() - 1
                plan = _DigestPlan()
                _prove_batch(plan, n_leaves, depth, positions)
                li = torch.tensor(plan.leaf_idx, dtype=torch.int64, device=self.device)
                ni = torch.tensor(plan.node_idx, dtype=torch.int64, device=self.device)
                lv = layer.leaves[li].cpu().numpy()
                nv = layer.nodes[ni].cpu().numpy() if plan.node_idx else np.empty((0, DIGEST_BYTES[self.hash_id]), dtype=np.uint8)
                proof = _prove_batch(_DigestTable(plan.leaf_idx, lv, plan.node_idx, nv), n_leaves, depth, positions)
                pi = torch.tensor(positions, dtype=torch.int64, device=self.device)
                vals = layer.evaluations[pi].cpu().numpy().view(np.uint64)
                if self.field == F64:
                    vals = vals.reshape(len(positions), -1)
                out.append(FriProofLayer(vals, proof))
                domain_size //= self.folding_factor()
        remainder = self.remainder_poly
        self.reset()
        return out, remainder
