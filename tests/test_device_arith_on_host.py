"""The kernels' field and hash code (csrc/field.cuh, csrc/hash.cuh) is __host__ __device__; this compiles it
with g++ (csrc/hosttest.cpp) and checks it against Python big ints, hashlib and the `blake3` module, so the
exact arithmetic the GPU runs is pinned without a GPU. Mirrors the reference's field tests
(math/src/field/f64/tests.rs:17-250, f128/tests.rs) and its hash call sites."""
import ctypes as C
import hashlib
import os
import random
import subprocess

import pytest

blake3 = pytest.importorskip("blake3")

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(os.path.dirname(HERE), "starkpack_winterfell_b200", "csrc")
P = 2**64 - 2**32 + 1
Q = 2**128 - 45 * 2**40 + 1
ROOT64 = 7277203076849721926
ROOT128 = 23953097886125630542083529559205016746
M64 = 2**64 - 1


@pytest.fixture(scope="module")
def L():
    so = os.path.join(CSRC, "libwg_hosttest.so")
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++",
                           os.path.join(CSRC, "hosttest.cpp"), "-o", so])
    lib = C.CDLL(so)
    lib.wgt_gl.restype = C.c_uint64
    lib.wgt_gl.argtypes = [C.c_int, C.c_uint64, C.c_uint64]
    return lib


EDGE64 = [0, 1, 2, P - 1, P - 2, 2**32 - 1, 2**32, 2**32 + 1, 2**63, P - 2**32, 0xFFFFFFFF00000000]


RINV = pow(2**64, -1, P)


def test_gl64_add_sub_mul(L):
    """words are Montgomery words; mul is the reference's Montgomery product (f64/mod.rs:356-361,651-661)."""
    rnd = random.Random(1)
    vals = EDGE64 + [rnd.randrange(P) for _ in range(1500)]
    for a in vals:
        for b in rnd.sample(vals, 4) + EDGE64:
            assert L.wgt_gl(0, a, b) == (a + b) % P
            assert L.wgt_gl(1, a, b) == (a - b) % P
            assert L.wgt_gl(2, a, b) == a * b * RINV % P
        assert L.wgt_gl(6, a, 0) == (-a) % P


def test_gl64_pow2_and_root16(L):
    rnd = random.Random(2)
    w16 = pow(ROOT64, 1 << 28, P)
    assert w16 == 2**12  # the order-64 root is 8 (f64/mod.rs:254-264)
    for k in range(192):
        assert L.wgt_gl(3, 0, k) == pow(2, k, P)
    for a in EDGE64 + [rnd.randrange(P) for _ in range(300)]:
        for i in range(16):  # linear map on the stored word, so it holds for Montgomery words as well
            assert L.wgt_gl(7, a, i) == a * pow(w16, i, P) % P
            assert L.wgt_gl(8, a, i) == a * pow(w16, -i, P) % P


def test_gl64_mont_conversions_and_roots(L):
    rnd = random.Random(3)
    for w in EDGE64 + [M64, P, P + 1] + [rnd.randrange(2**64) for _ in range(3000)]:
        assert L.wgt_gl(4, w, 0) == w * RINV % P  # as_int accepts any u64 (f64/mod.rs:275-282)
        assert L.wgt_gl(5, w, 0) == (w % P) * 2**64 % P
    for ln in range(33):  # math/src/field/traits.rs:254-263, returned as a Montgomery word
        assert L.wgt_gl(9, ln, 0) == pow(ROOT64, 1 << (32 - ln), P) * 2**64 % P
    assert L.wgt_gl(10, 12345 * 2**64 % P, 0) == pow(12345, -1, P) * 2**64 % P


def _f128(L, op, a, b):
    A = (C.c_uint64 * 2)(a & M64, a >> 64)
    B = (C.c_uint64 * 2)(b & M64, b >> 64)
    O = (C.c_uint64 * 2)()
    L.wgt_f128(op, A, B, O)
    return O[0] | (O[1] << 64)


def test_f128(L):
    rnd = random.Random(4)
    e = [0, 1, 2, Q - 1, Q - 2, 2**64 - 1, 2**64, 2**127, Q - 2**64, 2**128 - 2**46]
    vals = e + [rnd.randrange(Q) for _ in range(2000)] + [Q - 1 - rnd.randrange(2**50) for _ in range(200)]
    for a in vals:
        for b in rnd.sample(vals, 4) + e:
            assert _f128(L, 0, a, b) == (a + b) % Q
            assert _f128(L, 1, a, b) == (a - b) % Q
            assert _f128(L, 2, a, b) == (a * b) % Q
        assert _f128(L, 3, a, 0) == (-a) % Q
    w16 = pow(ROOT128, 1 << 36, Q)
    for a in vals[:150]:
        for i in range(16):
            assert _f128(L, 4, a, i) == a * pow(w16, i, Q) % Q
            assert _f128(L, 5, a, i) == a * pow(w16, -i, Q) % Q
    for ln in range(41):
        assert _f128(L, 6, ln, 0) == pow(ROOT128, 1 << (40 - ln), Q)
    assert _f128(L, 7, 987654321, 0) == pow(987654321, -1, Q)


def test_hashes_all_tree_shapes(L):
    rnd = random.Random(5)
    o = (C.c_uint8 * 32)()
    for n in list(range(0, 4300, 8)) + [1024 * 5 + 64, 1024 * 7, 8192, 65536, 65536 + 1024 + 8]:
        d = rnd.randbytes(n)
        L.wgt_hash(0, d, n, o)
        assert bytes(o) == blake3.blake3(d).digest(), n
        L.wgt_hash(1, d, n, o)
        assert bytes(o) == hashlib.sha3_256(d).digest(), n
    d = rnd.randbytes(64)
    L.wgt_b3_merge(d, o)
    assert bytes(o) == blake3.blake3(d).digest()
