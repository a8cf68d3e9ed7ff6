"""Device-resident front end: torch owns the HBM buffers and the stream, the C ABI does the work.
This is the boundary bench.py times for `value` (inputs already in HBM) and the piece a multi-GPU job runs
per rank (one process per GPU)."""
import ctypes as C

import torch

from . import _lib
from ._lib import ELEM_BYTES, F64, check, lib


def _vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class TraceCommitter:
    """Preallocates every buffer of Prover::build_trace_commitment (prover/src/lib.rs:615-670) for a fixed shape:
    cols/polys [sum(cols), n] elements, one row-major LDE [n*blowup, 8*ceil(cols_t/8)] per trace, leaves and
    nodes [n*blowup, 32] bytes, plus the NTT scratch."""

    def __init__(self, field, hash_id, n_rows, blowup, cols_per_trace, device="cuda:0", lde_scratch_cols=None):
        self.field, self.hash_id, self.n, self.blowup = field, hash_id, n_rows, blowup
        self.cols_per_trace = list(cols_per_trace)
        self.device = torch.device(device)
        _lib.init(self.device.index or 0)
        w = ELEM_BYTES[field] // 8  # int64 words per element
        tc = sum(self.cols_per_trace)
        L = n_rows * blowup
        kw = dict(dtype=torch.int64, device=self.device)
        self.cols = torch.empty((tc, n_rows, w), **kw)
        self.polys = torch.empty((tc, n_rows, w), **kw)
        self.row_widths = [(c + 7) // 8 * 8 for c in self.cols_per_trace]
        self.ldes = [torch.empty((L, rw, w), **kw) for rw in self.row_widths]
        self.leaves = torch.empty((L, 32), dtype=torch.uint8, device=self.device)
        self.nodes = torch.empty((L, 32), dtype=torch.uint8, device=self.device)
        self._cpt = (C.c_uint32 * len(self.cols_per_trace))(*self.cols_per_trace)
        if lde_scratch_cols is None:
            # the fastest plan's scratch (several narrow traces are extended together: wg_commit_scratch_elems)
            se = lib().wg_commit_scratch_elems(field, n_rows, blowup, len(self.cols_per_trace), self._cpt)
        else:
            se = max(lib().wg_interpolate_scratch_elems(field, tc, n_rows), lib().wg_lde_scratch_elems(field, lde_scratch_cols, n_rows, blowup))
        self.scratch_elems = int(se)
        self.scratch = torch.empty((max(self.scratch_elems, 1), w), **kw)
        self._cpt = (C.c_uint32 * len(self.cols_per_trace))(*self.cols_per_trace)
        self._lde_ptrs = (C.c_void_p * len(self.ldes))(*[t.data_ptr() for t in self.ldes])

    def commit(self, stream=None):
        """Runs the whole path on `stream` (default: torch's current stream); returns nothing — results stay in
        self.polys / self.ldes / self.leaves / self.nodes (root = self.nodes[1])."""
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        check(lib().wg_commit_trace_dev(self.field, self.hash_id, self.n, self.blowup, len(self.cols_per_trace), self._cpt,
                                        _vp(self.cols), _vp(self.polys), self._lde_ptrs, _vp(self.leaves), _vp(self.nodes),
                                        _vp(self.scratch), self.scratch_elems, C.c_void_p(s.cuda_stream)))

    def root(self):
        return bytes(self.nodes[1].cpu().numpy())
