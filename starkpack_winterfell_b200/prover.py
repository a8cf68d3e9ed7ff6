"""Mirror of Prover::build_trace_commitment (prover/src/lib.rs:615-670) and the containers it fills
(TraceLde trace_lde.rs:21-37, TraceCommitment commitment.rs:21-56, TracePolyTable poly_table.rs:21-37)."""
import ctypes as C

import numpy as np

from ._lib import DIGEST_BYTES, F64, F128, check, lib
from .matrix import ColMatrix, RowMatrix, StarkDomain, _ptr_array
from .merkle import MerkleTree, MerkleTreeError, _DigestPlan, _DigestTable, _prove, _prove_batch


class TraceCommitment:
    """prover/src/trace/commitment.rs:21-56."""

    def __init__(self, traces_lde, main_segment_tree, blowups):
        assert all(l.num_rows() == main_segment_tree.leaves().shape[0] for l in traces_lde), \
            "number of rows in trace LDE must be the same as number of leaves in trace commitment"
        self.traces_lde, self.main_segment_tree, self.blowups = traces_lde, main_segment_tree, blowups

    def root(self):
        return self.main_segment_tree.root()


class EvaluationFrame:
    """air/src/air/mod.rs EvaluationFrame: the current and the next row of the trace at one LDE step."""

    def __init__(self, num_columns, field=F64):
        shape = (num_columns,) if field == F64 else (num_columns, 2)
        self.current, self.next = np.zeros(shape, dtype=np.uint64), np.zeros(shape, dtype=np.uint64)


class TraceLde:
    """prover/src/trace/trace_lde.rs:21-108: owns the main-segment LDE (RowMatrix of base elements), the auxiliary
    segment LDEs and the blowup, and hands evaluation frames to the constraint evaluator. The main segment may also be a
    ResidentCommitment (the LDE stays in HBM and a frame is two gathered rows, wg_commitment_rows)."""

    def __init__(self, main_trace_lde, blowup, trace_index=0):
        self.main_segment_lde, self.aux_segment_ldes, self._blowup, self._trace = main_trace_lde, [], blowup, trace_index

    def _resident(self):
        return isinstance(self.main_segment_lde, ResidentCommitment)

    def add_aux_segment(self, aux_segment_lde):
        assert self.trace_len() == aux_segment_lde.num_rows(), \
            "number of rows in auxiliary segment must be of the same as in the main segment"
        self.aux_segment_ldes.append(aux_segment_lde)

    def main_trace_width(self):
        return self.main_segment_lde.cols_per_trace[self._trace] if self._resident() else self.main_segment_lde.num_cols()

    def aux_trace_width(self):
        return sum(m.num_cols() for m in self.aux_segment_ldes)

    def trace_len(self):
        return self.main_segment_lde.lde_rows if self._resident() else self.main_segment_lde.num_rows()

    def blowup(self):
        return self._blowup

    def read_main_trace_frame_into(self, lde_step, frame):
        # at the end of the trace, next state wraps around and we read the first step again (trace_lde.rs:78-90)
        nxt = (lde_step + self._blowup) % self.trace_len()
        if self._resident():
            rows = self.main_segment_lde.rows(self._trace, [lde_step, nxt])
            frame.current[...] = rows[0].reshape(frame.current.shape)
            frame.next[...] = rows[1].reshape(frame.next.shape)
        else:
            frame.current[...] = self.main_segment_lde.row(lde_step)
            frame.next[...] = self.main_segment_lde.row(nxt)

    def read_aux_trace_frame_into(self, lde_step, frame):
        nxt = (lde_step + self._blowup) % self.trace_len()
        seg = self.aux_segment_ldes[0]
        frame.current[...] = seg.row(lde_step)
        frame.next[...] = seg.row(nxt)

    def get_main_segment(self):
        return self.main_segment_lde

    def get_aux_segment(self, idx):
        return self.aux_segment_ldes[idx]


class TracePolyTable:
    """prover/src/trace/poly_table.rs:21-100: the trace polynomials in coefficient form (E = B here: out-of-domain points in
    an extension field belong to the DEEP composition, which stays with the caller)."""

    def __init__(self, main_trace_polys):
        self.main_segment_polys, self.aux_segment_polys = main_trace_polys, []

    def add_aux_segment(self, aux_segment_polys):
        assert self.poly_size() == aux_segment_polys.num_rows(), \
            "polynomials in auxiliary segment must be of the same size as in the main segment"
        self.aux_segment_polys.append(aux_segment_polys)

    def poly_size(self):
        return self.main_segment_polys.num_rows()

    def evaluate_at(self, x):
        """ColMatrix::evaluate_columns_at of every segment (col_matrix.rs:210-216), x an element in internal representation."""
        out = []
        for m in [self.main_segment_polys] + self.aux_segment_polys:
            f = m.field
            x = np.ascontiguousarray(x, dtype=np.uint64)
            res = np.empty((m.num_cols(),) if f == F64 else (m.num_cols(), 2), dtype=np.uint64)
            check(lib().wg_evaluate_columns_at(f, _ptr_array(m.columns), m.num_cols(), m.num_rows(), C.c_void_p(x.ctypes.data),
                                               C.c_void_p(res.ctypes.data)))
            out.append(res)
        return np.concatenate(out)

    def get_ood_frame(self, z, g):
        """[evaluate_at(z), evaluate_at(z * g)] (poly_table.rs:73-80); the caller supplies z * g (field products stay with it)."""
        return [self.evaluate_at(z), self.evaluate_at(g)]

    def main_trace_polys(self):
        return iter(self.main_segment_polys.columns)

    def num_main_trace_polys(self):
        return self.main_segment_polys.num_cols()

    def get_main_trace_poly(self, idx):
        return self.main_segment_polys.columns[idx]


def build_trace_commitment(hash_id, traces, domain, want_lde=True, want_polys=True):
    """traces: list of ColMatrix (one per aggregated trace, all with the same number of rows).
    Returns (list[RowMatrix], MerkleTree, list[ColMatrix]) like the reference."""
    f = traces[0].field
    deg = getattr(traces[0], "ext_degree", 1)  # E != B: auxiliary trace segments (prover/src/lib.rs:333-334)
    n, B = traces[0].num_rows(), domain.blowup
    assert all(t.num_rows() == n and t.field == f and getattr(t, "ext_degree", 1) == deg for t in traces)
    L = n * B
    cols = [c for t in traces for c in t.columns]
    cpt = (C.c_uint32 * len(traces))(*[t.num_cols() for t in traces])
    polys = [[np.empty_like(c) for c in t.columns] for t in traces] if want_polys else None
    ldes = None
    if want_lde:
        ldes = []
        for t in traces:
            rw = (t.num_cols() * deg + 7) // 8 * 8
            ldes.append(np.empty((L, rw) if f == F64 else (L, rw, 2), dtype=np.uint64))
    leaves = np.empty((L, DIGEST_BYTES[hash_id]), dtype=np.uint8)
    nodes = np.empty((L, DIGEST_BYTES[hash_id]), dtype=np.uint8)
    root = np.empty(DIGEST_BYTES[hash_id], dtype=np.uint8)
    args = (n, B, len(traces), cpt, _ptr_array(cols),
            _ptr_array([c for p in polys for c in p]) if polys else None,
            _ptr_array(ldes) if ldes else None,
            C.c_void_p(leaves.ctypes.data), C.c_void_p(nodes.ctypes.data), C.c_void_p(root.ctypes.data))
    if deg != 1:
        check(lib().wg_commit_trace_ext(f, hash_id, deg, *args))
    else:
        check(lib().wg_commit_trace(f, hash_id, *args))
    tree = MerkleTree.from_raw_parts(hash_id, nodes, leaves)
    lde_m = [RowMatrix(f, d, d.shape[1], t.num_cols() * deg) for d, t in zip(ldes, traces)] if ldes else []
    poly_m = [ColMatrix(f, p, deg) for p in polys] if polys else []
    return lde_m, tree, poly_m


def build_constraint_commitment(hash_id, composition_poly, domain):
    """Prover::build_constraint_commitment (prover/src/lib.rs:680-715): evaluate the composition polynomial columns
    (a ColMatrix of coefficients) over the LDE domain and commit to the rows. Returns (RowMatrix, MerkleTree), the two
    members of ConstraintCommitment (prover/src/constraints/commitment.rs)."""
    f = composition_poly.field
    n, B = composition_poly.num_rows(), domain.blowup
    L = n * B
    rw = (composition_poly.num_cols() + 7) // 8 * 8
    lde = np.empty((L, rw) if f == F64 else (L, rw, 2), dtype=np.uint64)
    leaves = np.empty((L, DIGEST_BYTES[hash_id]), dtype=np.uint8)
    nodes = np.empty((L, DIGEST_BYTES[hash_id]), dtype=np.uint8)
    root = np.empty(DIGEST_BYTES[hash_id], dtype=np.uint8)
    check(lib().wg_commit_polys(f, hash_id, n, B, composition_poly.num_cols(), _ptr_array(composition_poly.columns),
                                C.c_void_p(lde.ctypes.data), C.c_void_p(leaves.ctypes.data), C.c_void_p(nodes.ctypes.data),
                                C.c_void_p(root.ctypes.data)))
    return RowMatrix(f, lde, rw, composition_poly.num_cols()), MerkleTree.from_raw_parts(hash_id, nodes, leaves)


class JointTraceQueries:
    """prover/src/trace/commitment.rs build_segment_queries: the batch proof, the concatenated rows ("comb_states") and
    the per-trace rows at the queried positions."""

    def __init__(self, trace_proof, comb_states, traces_states):
        self.trace_proof, self.comb_states, self.traces_states = trace_proof, comb_states, traces_states


class ResidentCommitment:
    """A trace (or constraint) commitment whose LDE matrices and Merkle tree stay in HBM. `query(positions)` is
    TraceCommitment::query (prover/src/trace/commitment.rs:87-116): only the queried rows and the digests
    MerkleTree::prove_batch reads cross PCIe."""

    def __init__(self, field, hash_id, handle, root, polys):
        self.field, self.hash_id, self._h, self._root, self.polys = field, hash_id, handle, root, polys
        L, nt = C.c_uint64(), C.c_uint32()
        check(lib().wg_commitment_info(self._h, C.byref(L), C.byref(nt), None, None))
        cpt, rws = (C.c_uint32 * nt.value)(), (C.c_uint32 * nt.value)()
        check(lib().wg_commitment_info(self._h, None, None, cpt, rws))
        self.lde_rows, self.cols_per_trace, self.row_widths = L.value, list(cpt), list(rws)

    @classmethod
    def from_traces(cls, hash_id, traces, domain, want_polys=True, polys_out=None):
        """polys_out: optional preallocated arrays (e.g. pinned memory), one list per trace, that receive the polys."""
        f = traces[0].field
        n = traces[0].num_rows()
        assert all(t.num_rows() == n and t.field == f for t in traces)
        cols = [c for t in traces for c in t.columns]
        cpt = (C.c_uint32 * len(traces))(*[t.num_cols() for t in traces])
        polys = None
        if polys_out is not None:
            polys = polys_out
        elif want_polys:
            polys = [[np.empty_like(c) for c in t.columns] for t in traces]
        root = np.empty(DIGEST_BYTES[hash_id], dtype=np.uint8)
        h = C.c_void_p()
        check(lib().wg_commitment_from_trace(f, hash_id, n, domain.blowup, len(traces), cpt, _ptr_array(cols),
                                             _ptr_array([c for p in polys for c in p]) if polys else None,
                                             C.c_void_p(root.ctypes.data), C.byref(h)))
        return cls(f, hash_id, h, bytes(root), [ColMatrix(f, p) for p in polys] if polys else [])

    @classmethod
    def from_polys(cls, hash_id, composition_poly, domain):
        f = composition_poly.field
        root = np.empty(DIGEST_BYTES[hash_id], dtype=np.uint8)
        h = C.c_void_p()
        check(lib().wg_commitment_from_polys(f, hash_id, composition_poly.num_rows(), domain.blowup, composition_poly.num_cols(),
                                             _ptr_array(composition_poly.columns), C.c_void_p(root.ctypes.data), C.byref(h)))
        return cls(f, hash_id, h, bytes(root), [])

    def root(self):
        return self._root

    def depth(self):
        return self.lde_rows.bit_length() - 1

    def rows(self, trace, positions):
        """RowMatrix::row for each position: [n_pos, cols] (f64) / [n_pos, cols, 2] (f128)."""
        pos = np.ascontiguousarray(positions, dtype=np.uint64)
        c = self.cols_per_trace[trace]
        out = np.empty((len(pos), c) if self.field == F64 else (len(pos), c, 2), dtype=np.uint64)
        check(lib().wg_commitment_rows(self._h, trace, len(pos), C.c_void_p(pos.ctypes.data), C.c_void_p(out.ctypes.data)))
        return out

    def _digests(self, which, idx):
        idx = np.ascontiguousarray(idx, dtype=np.uint64)
        out = np.empty((len(idx), DIGEST_BYTES[self.hash_id]), dtype=np.uint8)
        if len(idx):
            check(lib().wg_commitment_digests(self._h, which, len(idx), C.c_void_p(idx.ctypes.data), C.c_void_p(out.ctypes.data)))
        return out

    def _with_digests(self, fn):
        plan = _DigestPlan()
        fn(plan)  # raises the reference's errors before touching the device
        table = _DigestTable(plan.leaf_idx, self._digests(0, plan.leaf_idx), plan.node_idx, self._digests(1, plan.node_idx))
        return fn(table)

    def prove(self, index):
        return self._with_digests(lambda src: _prove(src, self.lde_rows, int(index)))

    def prove_batch(self, indexes):
        return self._with_digests(lambda src: _prove_batch(src, self.lde_rows, self.depth(), indexes))

    def query(self, positions):
        traces_states = [self.rows(t, positions) for t in range(len(self.cols_per_trace))]
        comb_states = np.concatenate(traces_states, axis=1)
        return JointTraceQueries(self.prove_batch(positions), comb_states, traces_states)

    def download(self, want_lde=True):
        """-> (list[RowMatrix], MerkleTree): the full host copy build_trace_commitment returns."""
        L = self.lde_rows
        ldes = []
        if want_lde:
            for rw in self.row_widths:
                ldes.append(np.empty((L, rw) if self.field == F64 else (L, rw, 2), dtype=np.uint64))
        leaves = np.empty((L, DIGEST_BYTES[self.hash_id]), dtype=np.uint8)
        nodes = np.empty((L, DIGEST_BYTES[self.hash_id]), dtype=np.uint8)
        check(lib().wg_commitment_download(self._h, _ptr_array(ldes) if ldes else None, C.c_void_p(leaves.ctypes.data),
                                           C.c_void_p(nodes.ctypes.data)))
        mats = [RowMatrix(self.field, d, d.shape[1], c) for d, c in zip(ldes, self.cols_per_trace)]
        return mats, MerkleTree.from_raw_parts(self.hash_id, nodes, leaves)

    def close(self):
        if self._h:
            lib().wg_commitment_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
