"""Mirror of Prover::build_trace_commitment (prover/src/lib.rs:615-670) and the containers it fills
(TraceLde trace_lde.rs:21-37, TraceCommitment commitment.rs:21-56, TracePolyTable poly_table.rs:21-37)."""
import ctypes as C

import numpy as np

from ._lib import F64, F128, check, lib
from .matrix import ColMatrix, RowMatrix, StarkDomain, _ptr_array
from .merkle import MerkleTree


class TraceCommitment:
    """prover/src/trace/commitment.rs:21-56."""

    def __init__(self, traces_lde, main_segment_tree, blowups):
        assert all(l.num_rows() == main_segment_tree.leaves().shape[0] for l in traces_lde), \
            "number of rows in trace LDE must be the same as number of leaves in trace commitment"
        self.traces_lde, self.main_segment_tree, self.blowups = traces_lde, main_segment_tree, blowups

    def root(self):
        return self.main_segment_tree.root()


def build_trace_commitment(hash_id, traces, domain, want_lde=True, want_polys=True):
    """traces: list of ColMatrix (one per aggregated trace, all with the same number of rows).
    Returns (list[RowMatrix], MerkleTree, list[ColMatrix]) like the reference."""
    f = traces[0].field
    n, B = traces[0].num_rows(), domain.blowup
    assert all(t.num_rows() == n and t.field == f for t in traces)
    L = n * B
    cols = [c for t in traces for c in t.columns]
    cpt = (C.c_uint32 * len(traces))(*[t.num_cols() for t in traces])
    polys = [[np.empty_like(c) for c in t.columns] for t in traces] if want_polys else None
    ldes = None
    if want_lde:
        ldes = []
        for t in traces:
            rw = (t.num_cols() + 7) // 8 * 8
            ldes.append(np.empty((L, rw) if f == F64 else (L, rw, 2), dtype=np.uint64))
    leaves = np.empty((L, 32), dtype=np.uint8)
    nodes = np.empty((L, 32), dtype=np.uint8)
    root = np.empty(32, dtype=np.uint8)
    check(lib().wg_commit_trace(
        f, hash_id, n, B, len(traces), cpt, _ptr_array(cols),
        _ptr_array([c for p in polys for c in p]) if polys else None,
        _ptr_array(ldes) if ldes else None,
        C.c_void_p(leaves.ctypes.data), C.c_void_p(nodes.ctypes.data), C.c_void_p(root.ctypes.data)))
    tree = MerkleTree.from_raw_parts(hash_id, nodes, leaves)
    lde_m = [RowMatrix(f, d, d.shape[1], t.num_cols()) for d, t in zip(ldes, traces)] if ldes else []
    poly_m = [ColMatrix(f, p) for p in polys] if polys else []
    return lde_m, tree, poly_m
