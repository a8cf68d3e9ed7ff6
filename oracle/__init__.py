"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the C restatement of the reference's CPU path, oracle.c) plus a
pure-Python big-int ground truth (`pyref`) used to pin the C restatement on small cases.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this package. The product (starkpack_winterfell_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

DIGEST = {0: 32, 1: 32, 2: 24}  # digest bytes of Blake3_256, Sha3_256, Blake3_192

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

F64, F128 = 0, 1
BLAKE3_256, SHA3_256 = 0, 1
ESIZE = {F64: 8, F128: 16}


def build():
    """Compile oracle/liboracle.so with the committed Makefile (gcc -O3 -fopenmp)."""
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        L = _LIB
        for name in ("or_f64_new", "or_f64_as_int", "or_f64_inv"):
            getattr(L, name).restype = C.c_uint64
            getattr(L, name).argtypes = [C.c_uint64]
        for name in ("or_f64_add", "or_f64_sub", "or_f64_mul", "or_f64_exp"):
            getattr(L, name).restype = C.c_uint64
            getattr(L, name).argtypes = [C.c_uint64, C.c_uint64]
        L.or_threads.restype = C.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def elems(field, n):
    """uninitialised element buffer: f64 -> uint64[n]; f128 -> uint64[n,2] (lo, hi little-endian)."""
    return np.empty(n, dtype=np.uint64) if field == F64 else np.empty((n, 2), dtype=np.uint64)


def get_twiddles(field, n, inverse=False):
    out = elems(field, max(n // 2, 1))
    rc = lib().or_get_twiddles(field, C.c_uint64(n), int(inverse), _p(out))
    if rc:
        raise ValueError(f"or_get_twiddles rc={rc}")
    return out[: n // 2]


def interpolate_poly(field, evals, inv_tw=None):
    v = np.ascontiguousarray(evals).copy()
    n = v.shape[0]
    if inv_tw is None:
        inv_tw = get_twiddles(field, n, True)
    rc = lib().or_interpolate_poly(field, _p(v), C.c_uint64(n), _p(np.ascontiguousarray(inv_tw)))
    if rc:
        raise ValueError(f"or_interpolate_poly rc={rc}")
    return v


def evaluate_poly_with_offset(field, p, offset, blowup, tw=None):
    p = np.ascontiguousarray(p)
    n = p.shape[0]
    if tw is None:
        tw = get_twiddles(field, n)
    out = elems(field, n * blowup)
    off = np.ascontiguousarray(offset, dtype=np.uint64)
    rc = lib().or_evaluate_poly_with_offset(field, _p(p), C.c_uint64(n), _p(np.ascontiguousarray(tw)), _p(off),
                                            C.c_uint64(blowup), _p(out))
    if rc:
        raise ValueError(f"or_evaluate_poly_with_offset rc={rc}")
    return out


def hash_bytes(hash_id, data: bytes) -> bytes:
    out = (C.c_uint8 * 32)()
    lib().or_hash_bytes(hash_id, data, C.c_uint64(len(data)), out)
    return bytes(out)[: DIGEST[hash_id]]


def hash_rows(field, hash_id, row_major, rows, row_width, elems_per_row):
    leaves = np.empty((rows, DIGEST[hash_id]), dtype=np.uint8)
    rc = lib().or_hash_rows(field, hash_id, _p(np.ascontiguousarray(row_major)), C.c_uint64(rows), C.c_uint64(row_width),
                            C.c_uint64(elems_per_row), _p(leaves))
    assert rc == 0
    return leaves


def merkle_nodes(hash_id, leaves):
    leaves = np.ascontiguousarray(leaves, dtype=np.uint8).reshape(-1, DIGEST[hash_id])
    n = leaves.shape[0]
    nodes = np.empty((max(n, 1), DIGEST[hash_id]), dtype=np.uint8)
    rc = lib().or_merkle_nodes(hash_id, _p(leaves), C.c_uint64(n), _p(nodes))
    if rc:
        raise ValueError(f"or_merkle_nodes rc={rc}")
    return nodes


def commit_trace(field, hash_id, traces, blowup, want=("polys", "lde", "leaves", "nodes")):
    """traces: list of arrays [cols, n] (f64) or [cols, n, 2] (f128), column-major like ColMatrix.
    Returns dict(polys=[...], lde=[...] row-major [L, row_width(,2)], leaves, nodes, root)."""
    traces = [np.ascontiguousarray(t, dtype=np.uint64) for t in traces]
    n = traces[0].shape[1]
    L = n * blowup
    cols = [t.shape[0] for t in traces]
    ncols = sum(cols)
    col_ptrs = (C.c_void_p * ncols)()
    k = 0
    for t in traces:
        for c in range(t.shape[0]):
            col_ptrs[k] = t[c].ctypes.data
            k += 1
    polys = [np.empty_like(t) for t in traces] if "polys" in want else None
    poly_ptrs = None
    if polys is not None:
        poly_ptrs = (C.c_void_p * ncols)()
        k = 0
        for t in polys:
            for c in range(t.shape[0]):
                poly_ptrs[k] = t[c].ctypes.data
                k += 1
    ldes, lde_ptrs = None, None
    if "lde" in want:
        ldes = []
        lde_ptrs = (C.c_void_p * len(traces))()
        for i, t in enumerate(traces):
            rw = (t.shape[0] + 7) // 8 * 8
            shp = (L, rw) if field == F64 else (L, rw, 2)
            ldes.append(np.empty(shp, dtype=np.uint64))
            lde_ptrs[i] = ldes[-1].ctypes.data
    leaves = np.empty((L, DIGEST[hash_id]), dtype=np.uint8) if "leaves" in want else None
    nodes = np.empty((L, DIGEST[hash_id]), dtype=np.uint8) if "nodes" in want else None
    root = np.empty(DIGEST[hash_id], dtype=np.uint8)
    cpt = (C.c_uint32 * len(cols))(*cols)
    rc = lib().or_commit_trace(field, hash_id, C.c_uint64(n), C.c_uint64(blowup), C.c_uint32(len(traces)), cpt, col_ptrs,
                               poly_ptrs, lde_ptrs, _p(leaves), _p(nodes), _p(root))
    if rc:
        raise ValueError(f"or_commit_trace rc={rc}")
    return dict(polys=polys, lde=ldes, leaves=leaves, nodes=nodes, root=root)


def hash_comb_rows(field, hash_id, mats, elems):
    """commit_to_comb_rows leaves (row_matrix.rs:204-238) over several row-major matrices [rows, rw_i(,2)]."""
    mats = [np.ascontiguousarray(m, dtype=np.uint64) for m in mats]
    rows = mats[0].shape[0]
    k = len(mats)
    ptrs = (C.c_void_p * k)(*[m.ctypes.data for m in mats])
    rws = (C.c_uint64 * k)(*[m.shape[1] for m in mats])
    els = (C.c_uint64 * k)(*elems)
    leaves = np.empty((rows, DIGEST[hash_id]), dtype=np.uint8)
    rc = lib().or_hash_comb_rows(field, hash_id, C.c_uint32(k), ptrs, rws, els, C.c_uint64(rows), _p(leaves))
    assert rc == 0
    return leaves


def fri_commit_layer(field, hash_id, evals, folding):
    """FriProver::build_layer first half (fri/src/prover/mod.rs:185-199) -> dict(transposed, leaves, nodes, root)."""
    e = np.ascontiguousarray(evals, dtype=np.uint64)
    n = e.shape[0]
    rows = n // folding
    tr = np.empty_like(e)
    leaves = np.empty((rows, DIGEST[hash_id]), dtype=np.uint8)
    nodes = np.empty((rows, DIGEST[hash_id]), dtype=np.uint8)
    rc = lib().or_fri_commit_layer(field, hash_id, _p(e), C.c_uint64(n), C.c_uint32(folding), _p(tr), _p(leaves), _p(nodes))
    if rc:
        raise ValueError(f"or_fri_commit_layer rc={rc}")
    shape = (rows, folding) if field == F64 else (rows, folding, 2)
    return dict(transposed=tr.reshape(shape), leaves=leaves, nodes=nodes, root=bytes(nodes[1]))


def fri_apply_drp(field, transposed, folding, offset, alpha):
    """folding::apply_drp (fri/src/folding/mod.rs:115-146); transposed [rows, folding(,2)] -> [rows(,2)]."""
    t = np.ascontiguousarray(transposed, dtype=np.uint64)
    rows = t.shape[0]
    out = elems(field, rows)
    rc = lib().or_fri_apply_drp(field, _p(t), C.c_uint64(rows), C.c_uint32(folding),
                                _p(np.ascontiguousarray(offset, dtype=np.uint64)), _p(np.ascontiguousarray(alpha, dtype=np.uint64)), _p(out))
    if rc:
        raise ValueError(f"or_fri_apply_drp rc={rc}")
    return out


def interpolate_poly_ext(field, deg, evals, offset=None):
    """fft::interpolate_poly / interpolate_poly_with_offset for E = degree-`deg` extension; evals [n, deg(,2)]."""
    v = np.ascontiguousarray(evals, dtype=np.uint64).copy()
    n = v.shape[0]
    inv_tw = np.ascontiguousarray(get_twiddles(field, n, True))
    off = np.ascontiguousarray(offset, dtype=np.uint64) if offset is not None else None
    rc = lib().or_interpolate_poly_ext(field, C.c_uint32(deg), _p(v), C.c_uint64(n), _p(inv_tw), _p(off))
    if rc:
        raise ValueError(f"or_interpolate_poly_ext rc={rc}")
    return v


def evaluate_poly_with_offset_ext(field, deg, p, offset, blowup):
    p = np.ascontiguousarray(p, dtype=np.uint64)
    n = p.shape[0]
    tw = np.ascontiguousarray(get_twiddles(field, n))
    out = np.empty((n * blowup,) + p.shape[1:], dtype=np.uint64)
    rc = lib().or_evaluate_poly_with_offset_ext(field, C.c_uint32(deg), _p(p), C.c_uint64(n), _p(tw),
                                                _p(np.ascontiguousarray(offset, dtype=np.uint64)), C.c_uint64(blowup), _p(out))
    if rc:
        raise ValueError(f"or_evaluate_poly_with_offset_ext rc={rc}")
    return out


def split_ext_columns(field, cols):
    """[C, n, deg(,2)] extension columns -> [C*deg, n(,2)] base columns, base column c*deg + d = coordinate d of column c
    (ColMatrix::get_base_element, prover/src/matrix/col_matrix.rs:113-119)."""
    a = np.asarray(cols, dtype=np.uint64)
    a = np.moveaxis(a, 2, 1)  # [C, deg, n(,2)]
    return np.ascontiguousarray(a.reshape((a.shape[0] * a.shape[1],) + a.shape[2:]))


def merge_ext_columns(field, base_cols, deg):
    a = np.asarray(base_cols, dtype=np.uint64)
    a = a.reshape((a.shape[0] // deg, deg) + a.shape[1:])
    return np.ascontiguousarray(np.moveaxis(a, 1, 2))
