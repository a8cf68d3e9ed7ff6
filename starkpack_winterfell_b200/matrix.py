"""Mirror of the prover's matrix types (prover/src/matrix/{col_matrix,row_matrix}.rs) and StarkDomain
(prover/src/domain.rs) over the C ABI. Host numpy in, host numpy out; all arithmetic happens on the GPU."""
import ctypes as C

import numpy as np

from . import fft
from ._lib import DIGEST_BYTES, F64, F128, check, lib
from .merkle import MerkleTree

GENERATOR_INTERNAL = {  # GENERATOR in internal representation: f64 Montgomery word of 7; f128 canonical 3
    F64: np.array([(7 << 64) % (2**64 - 2**32 + 1)], dtype=np.uint64),
    F128: np.array([3, 0], dtype=np.uint64),
}


def _ptr_array(arrs):
    pa = (C.c_void_p * len(arrs))()
    for i, a in enumerate(arrs):
        pa[i] = a.ctypes.data
    return pa


class StarkDomain:
    """prover/src/domain.rs:13-77: trace twiddles, blowup, offset (= GENERATOR, air/src/options.rs:199-201)."""

    def __init__(self, field, trace_length, blowup, offset=None):
        self.field, self.trace_length, self.blowup = field, trace_length, blowup
        self.offset = GENERATOR_INTERNAL[field] if offset is None else np.ascontiguousarray(offset, dtype=np.uint64).reshape(-1)
        self._twiddles = None

    @classmethod
    def from_twiddles(cls, field, trace_twiddles, blowup, offset):
        d = cls(field, len(trace_twiddles) * 2, blowup, offset)
        d._twiddles = trace_twiddles
        return d

    def trace_twiddles(self):
        if self._twiddles is None:
            self._twiddles = fft.get_twiddles(self.field, self.trace_length)
        return self._twiddles

    def lde_domain_size(self):
        return self.trace_length * self.blowup


class ColMatrix:
    """prover/src/matrix/col_matrix.rs:31-69: a list of separately allocated columns."""

    def __init__(self, field, columns, ext_degree=1):
        """ext_degree > 1: ColMatrix<E> over a quadratic / cubic extension; a column is [n, deg] (f64) or [n, deg, 2]."""
        self.field = field
        self.ext_degree = ext_degree
        self.columns = [np.ascontiguousarray(c, dtype=np.uint64) for c in columns]
        assert len(self.columns) > 0, "a matrix must contain at least one column"
        n = self.columns[0].shape[0]
        assert n > 1 and n & (n - 1) == 0, "number of rows must be a power of two greater than 1"
        assert all(c.shape == self.columns[0].shape for c in self.columns), "all columns must have the same length"

    def num_cols(self):
        return len(self.columns)

    def num_rows(self):
        return self.columns[0].shape[0]

    def num_base_cols(self):
        """col_matrix.rs:78-80"""
        return len(self.columns) * self.ext_degree

    def interpolate_columns(self):
        """ColMatrix::interpolate_columns (col_matrix.rs:196-206): new matrix, input untouched."""
        outs = [np.empty_like(c) for c in self.columns]
        if self.ext_degree != 1:
            check(lib().wg_interpolate_columns_ext(self.field, self.ext_degree, _ptr_array(self.columns), len(outs), self.num_rows(),
                                                   _ptr_array(outs)))
            return ColMatrix(self.field, outs, self.ext_degree)
        check(lib().wg_interpolate_columns(self.field, _ptr_array(self.columns), len(outs), self.num_rows(), _ptr_array(outs)))
        return ColMatrix(self.field, outs)


class RowMatrix:
    """prover/src/matrix/row_matrix.rs:31-39: row-major data, row_width padded to 8, elements_per_row real."""

    def __init__(self, field, data, row_width, elements_per_row):
        self.field, self.data, self.row_width, self.elements_per_row = field, data, row_width, elements_per_row

    @classmethod
    def evaluate_polys_over(cls, polys, domain):
        """RowMatrix::evaluate_polys_over::<8> (row_matrix.rs:82-98)."""
        f, n, B, C_ = polys.field, polys.num_rows(), domain.blowup, polys.num_cols()
        deg = getattr(polys, "ext_degree", 1)
        if deg != 1:  # RowMatrix<E> stores base elements: num_base_cols per row (row_matrix.rs:33-41,97)
            nb = C_ * deg
            rw = (nb + 7) // 8 * 8
            out = np.empty((n * B, rw) if f == F64 else (n * B, rw, 2), dtype=np.uint64)
            check(lib().wg_evaluate_polys_over_ext(f, deg, _ptr_array(polys.columns), C_, n, B, C.c_void_p(domain.offset.ctypes.data),
                                                   C.c_void_p(out.ctypes.data)))
            return cls(f, out, rw, nb)
        rw = (C_ + 7) // 8 * 8
        shape = (n * B, rw) if f == F64 else (n * B, rw, 2)
        out = np.empty(shape, dtype=np.uint64)
        check(lib().wg_evaluate_polys_over(f, _ptr_array(polys.columns), C_, n, B, C.c_void_p(domain.offset.ctypes.data),
                                           C.c_void_p(out.ctypes.data)))
        return cls(f, out, rw, C_)

    @classmethod
    def evaluate_polys(cls, polys, blowup):
        """RowMatrix::evaluate_polys::<8> (row_matrix.rs:55-71)."""
        return cls.evaluate_polys_over(polys, StarkDomain(polys.field, polys.num_rows(), blowup))

    def num_rows(self):
        return self.data.shape[0]

    def num_cols(self):
        return self.elements_per_row

    def row(self, i):
        return self.data[i, : self.elements_per_row]

    def hash_rows(self, hash_id):
        leaves = np.empty((self.num_rows(), DIGEST_BYTES[hash_id]), dtype=np.uint8)
        check(lib().wg_hash_rows(self.field, hash_id, C.c_void_p(self.data.ctypes.data), self.num_rows(), self.row_width,
                                 self.elements_per_row, C.c_void_p(leaves.ctypes.data)))
        return leaves

    def commit_to_rows(self, hash_id):
        """RowMatrix::commit_to_rows (row_matrix.rs:183-203)."""
        return MerkleTree.new(hash_id, self.hash_rows(hash_id))

    def commit_to_comb_rows(self, hash_id, traces_lde):
        """RowMatrix::commit_to_comb_rows (row_matrix.rs:204-238): one tree whose leaf k hashes row k of every matrix in
        `traces_lde`, concatenated; as in the reference `self` must be traces_lde[0] (it skips index 0, :220)."""
        assert traces_lde and traces_lde[0] is self, "self must be the first of traces_lde (row_matrix.rs:220)"
        rows = self.num_rows()
        assert all(m.num_rows() == rows and m.field == self.field for m in traces_lde)
        k = len(traces_lde)
        ptrs = (C.c_void_p * k)(*[m.data.ctypes.data for m in traces_lde])
        rws = (C.c_uint64 * k)(*[m.row_width for m in traces_lde])
        els = (C.c_uint64 * k)(*[m.elements_per_row for m in traces_lde])
        leaves = np.empty((rows, DIGEST_BYTES[hash_id]), dtype=np.uint8)
        check(lib().wg_hash_comb_rows(self.field, hash_id, k, ptrs, rws, els, rows, C.c_void_p(leaves.ctypes.data)))
        return MerkleTree.new(hash_id, leaves)
