"""Mirror of winter_math::fft (math/src/fft/mod.rs) over the C ABI. Arrays are numpy uint64: f64 -> shape [n]
(Montgomery words, the reference's internal representation), f128 -> shape [n, 2] (lo, hi).

The reference's functions are generic over E: FieldElement<BaseField = B>; `ext_degree` (1, 2, 3) selects E = B,
QuadExtension<B>, CubeExtension<B>: an element is ext_degree consecutive base elements, arrays get one more axis
(f64 [n, deg], f128 [n, deg, 2]); twiddles and offsets stay base-field elements."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import F64, F128, check, lib


def _elems(field, n):
    return np.empty(n, dtype=np.uint64) if field == F64 else np.empty((n, 2), dtype=np.uint64)


def _p(a):
    return C.c_void_p(a.ctypes.data)


def _n(field, a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if field == F128 and (a.ndim != 2 or a.shape[1] != 2):
        raise ValueError("f128 arrays have shape [n, 2]")
    return a


def get_twiddles(field, domain_size):
    """fft::get_twiddles (math/src/fft/mod.rs:466-482)."""
    out = _elems(field, max(domain_size // 2, 1))
    check(lib().wg_get_twiddles(field, domain_size, 0, _p(out)))
    return out[: domain_size // 2]


def get_inv_twiddles(field, domain_size):
    """fft::get_inv_twiddles (math/src/fft/mod.rs:505-522)."""
    out = _elems(field, max(domain_size // 2, 1))
    check(lib().wg_get_twiddles(field, domain_size, 1, _p(out)))
    return out[: domain_size // 2]


def interpolate_poly(field, evaluations, inv_twiddles=None, ext_degree=1):
    """fft::interpolate_poly (math/src/fft/mod.rs:274-305): IN PLACE on `evaluations` (must be a contiguous
    uint64 numpy array). Same preconditions as the reference: power-of-two length, twiddles of length n/2."""
    v = evaluations
    if not (isinstance(v, np.ndarray) and v.dtype == np.uint64 and v.flags.c_contiguous):
        raise ValueError("evaluations must be a contiguous uint64 numpy array (modified in place)")
    n = v.shape[0]
    if inv_twiddles is not None and len(inv_twiddles) * 2 != n:
        raise ValueError(f"invalid number of twiddles: expected {n // 2} but received {len(inv_twiddles)}")
    if ext_degree != 1:
        return check(lib().wg_interpolate_poly_ext(field, ext_degree, _p(v), n, None))
    check(lib().wg_interpolate_poly(field, _p(v), n, None))


def evaluate_poly_with_offset(field, p, twiddles, domain_offset, blowup_factor, ext_degree=1):
    """fft::evaluate_poly_with_offset (math/src/fft/mod.rs:171-220). domain_offset is one element in internal
    representation (uint64 for f64 = Montgomery word; [lo, hi] for f128)."""
    p = _n(field, p) if ext_degree == 1 else np.ascontiguousarray(p, dtype=np.uint64)
    n = p.shape[0]
    if twiddles is not None and len(twiddles) * 2 != n:
        raise ValueError(f"invalid number of twiddles: expected {n // 2} but received {len(twiddles)}")
    off = np.ascontiguousarray(domain_offset, dtype=np.uint64).reshape(-1)
    if ext_degree != 1:
        out = np.empty((n * blowup_factor,) + p.shape[1:], dtype=np.uint64)
        check(lib().wg_evaluate_poly_with_offset_ext(field, ext_degree, _p(p), n, None, _p(off), blowup_factor, _p(out)))
        return out
    out = _elems(field, n * blowup_factor)
    check(lib().wg_evaluate_poly_with_offset(field, _p(p), n, None, _p(off), blowup_factor, _p(out)))
    return out


def _inplace(v, what):
    if not (isinstance(v, np.ndarray) and v.dtype == np.uint64 and v.flags.c_contiguous):
        raise ValueError(f"{what} must be a contiguous uint64 numpy array (modified in place)")
    return v.shape[0]


def evaluate_poly(field, p, twiddles=None, ext_degree=1):
    """fft::evaluate_poly (math/src/fft/mod.rs:85-116): IN PLACE, p[k] <- P(w_n^k), natural order."""
    n = _inplace(p, "p")
    if twiddles is not None and len(twiddles) * 2 != n:
        raise ValueError(f"invalid number of twiddles: expected {n // 2} but received {len(twiddles)}")
    if ext_degree != 1:
        return check(lib().wg_evaluate_poly_ext(field, ext_degree, _p(p), n, None))
    check(lib().wg_evaluate_poly(field, _p(p), n, None))


def serial_fft(field, values, twiddles=None, ext_degree=1):
    """fft::serial_fft (math/src/fft/mod.rs:416-440): fft_in_place followed by permute, i.e. the natural-order transform."""
    evaluate_poly(field, values, twiddles, ext_degree)


def interpolate_poly_with_offset(field, evaluations, inv_twiddles, domain_offset, ext_degree=1):
    """fft::interpolate_poly_with_offset (math/src/fft/mod.rs:362-400): IN PLACE; evaluations were taken over the coset
    domain_offset * <w_n>. domain_offset is one element in internal representation."""
    n = _inplace(evaluations, "evaluations")
    if inv_twiddles is not None and len(inv_twiddles) * 2 != n:
        raise ValueError(f"invalid number of twiddles: expected {n // 2} but received {len(inv_twiddles)}")
    off = np.ascontiguousarray(domain_offset, dtype=np.uint64).reshape(-1)
    if ext_degree != 1:
        return check(lib().wg_interpolate_poly_with_offset_ext(field, ext_degree, _p(evaluations), n, None, _p(off)))
    check(lib().wg_interpolate_poly_with_offset(field, _p(evaluations), n, None, _p(off)))


def infer_degree(field, evaluations, domain_offset):
    """fft::infer_degree (math/src/fft/mod.rs:540-560): degree of the polynomial interpolating `evaluations` over the
    coset; 0 for the zero polynomial (polynom::degree_of)."""
    poly = np.array(_n(field, evaluations), copy=True)
    interpolate_poly_with_offset(field, poly, None, domain_offset)
    nz = np.nonzero(poly.reshape(poly.shape[0], -1).any(axis=1))[0]
    return int(nz[-1]) if len(nz) else 0


def permute_index(size, index):
    """fft::permute_index (math/src/fft/mod.rs:592-600)."""
    bits = size.bit_length() - 1
    return int(format(index, f"0{bits}b")[::-1], 2) if bits else 0
