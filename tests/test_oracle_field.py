"""Pins oracle.c field arithmetic to Python big ints + the reference's own constants
(math/src/field/f64/tests.rs:17-160, f128/tests.rs:129-141)."""
import ctypes as C
import random

import numpy as np

from oracle import pyref

P64, P128 = pyref.P64, pyref.P128
R = 2**64 % P64


def test_f64_ops_vs_bigint(oracle):
    L = oracle.lib()
    rnd = random.Random(1)
    edge = [0, 1, 2, P64 - 1, P64 - 2, 2**32, 2**32 - 1, 2**63, 0xFFFFFFFF00000000]
    vals = edge + [rnd.randrange(P64) for _ in range(300)]
    for a in vals:
        ma = L.or_f64_new(a)
        assert ma == a * R % P64 and ma < P64
        assert L.or_f64_as_int(ma) == a
        for b in vals[:40]:
            mb = L.or_f64_new(b)
            assert L.or_f64_as_int(L.or_f64_add(ma, mb)) == (a + b) % P64
            assert L.or_f64_as_int(L.or_f64_sub(ma, mb)) == (a - b) % P64
            assert L.or_f64_as_int(L.or_f64_mul(ma, mb)) == (a * b) % P64
        if a:
            assert L.or_f64_as_int(L.or_f64_mul(ma, L.or_f64_inv(ma))) == 1
    # new() reduces values >= p (f64/mod.rs:57-59)
    assert L.or_f64_as_int(L.or_f64_new(P64 + 5)) == 5


def _f128(L, op, a, b=0):
    x = np.array([a & (2**64 - 1), a >> 64], dtype=np.uint64)
    y = np.array([b & (2**64 - 1), b >> 64], dtype=np.uint64)
    r = np.zeros(2, dtype=np.uint64)
    L.or_f128_op(op, x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), r.ctypes.data_as(C.c_void_p))
    return int(r[0]) | (int(r[1]) << 64)


def test_f128_ops_vs_bigint(oracle):
    L = oracle.lib()
    rnd = random.Random(2)
    edge = [0, 1, 2, P128 - 1, P128 - 2, 2**64, 2**64 - 1, 2**127, 45 * 2**40, 2**128 - 45 * 2**40]
    vals = edge + [rnd.randrange(P128) for _ in range(200)]
    for a in vals:
        for b in vals[:40]:
            assert _f128(L, 0, a, b) == (a + b) % P128
            assert _f128(L, 1, a, b) == (a - b) % P128
            assert _f128(L, 2, a, b) == (a * b) % P128
        if a:
            assert _f128(L, 2, a, _f128(L, 3, a)) == 1


def test_roots_of_unity(oracle):
    """f64/tests.rs:150-160, f128/tests.rs:129-141: root^(2^adicity) == 1, order exact; w64 = 8 in f64."""
    L = oracle.lib()
    for field, p in ((0, P64), (1, P128)):
        for log_n in (1, 2, 6, 10, 20, pyref.TWO_ADICITY[field]):
            out = np.zeros(2, dtype=np.uint64)
            L.or_root_of_unity(field, log_n, out.ctypes.data_as(C.c_void_p))
            got = pyref.array_to_ints(field, out[:1] if field == 0 else out)[0]
            assert got == pyref.root_of_unity(field, log_n)
            assert pow(got, 1 << log_n, p) == 1 and pow(got, 1 << (log_n - 1), p) == p - 1
    assert pyref.root_of_unity(0, 6) == 8
    assert pow(3, (P128 - 1) >> 40, P128) == pyref.ROOT[1]
