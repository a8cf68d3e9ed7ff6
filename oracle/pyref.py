"""ORACLE — TEST INFRASTRUCTURE ONLY. Pure-Python big-int ground truth for SMALL cases.

This is the definition the reference's own tests check its FFT against: O(n^2) `polynom::eval_many`
over an explicit (shifted) domain (math/src/fft/tests.rs:19-60, prover/src/matrix/tests.rs:14-40,
doc-tests math/src/fft/mod.rs:147-170), plus hashing via the independent Python `blake3` module and
hashlib.sha3_256. It pins oracle.c, which in turn checks the CUDA path at larger sizes.
"""
import hashlib

import numpy as np

try:
    import blake3 as _blake3
except ImportError:  # pragma: no cover
    _blake3 = None

P64 = 2**64 - 2**32 + 1
P128 = 2**128 - 45 * 2**40 + 1
MOD = {0: P64, 1: P128}
TWO_ADICITY = {0: 32, 1: 40}
ROOT = {0: 7277203076849721926, 1: 23953097886125630542083529559205016746}  # f64/mod.rs:264, f128/mod.rs:38
GENERATOR = {0: 7, 1: 3}  # f64/mod.rs:248, f128/mod.rs:165
R64 = 2**64 % P64


def root_of_unity(field, log_n):
    """math/src/field/traits.rs:254-263 (canonical integer)."""
    return pow(ROOT[field], 1 << (TWO_ADICITY[field] - log_n), MOD[field])


def to_internal(field, v):
    """canonical int -> stored word (f64: Montgomery x*2^64 mod p; f128: canonical)."""
    return (v * R64) % P64 if field == 0 else v


def from_internal(field, w):
    return (w * pow(R64, -1, P64)) % P64 if field == 0 else w


def ints_to_array(field, vals):
    """canonical ints -> numpy array in the stored representation."""
    if field == 0:
        return np.array([to_internal(0, v) for v in vals], dtype=np.uint64)
    a = np.empty((len(vals), 2), dtype=np.uint64)
    for i, v in enumerate(vals):
        a[i, 0] = v & (2**64 - 1)
        a[i, 1] = v >> 64
    return a


def array_to_ints(field, arr):
    """numpy stored representation -> canonical ints."""
    arr = np.asarray(arr)
    if field == 0:
        return [from_internal(0, int(w)) for w in arr.reshape(-1)]
    a = arr.reshape(-1, 2)
    return [int(a[i, 0]) | (int(a[i, 1]) << 64) for i in range(a.shape[0])]


def bitrev(i, bits):
    r = 0
    for b in range(bits):
        r |= ((i >> b) & 1) << (bits - 1 - b)
    return r


def twiddles(field, n, inverse=False):
    """math/src/fft/mod.rs:466-522: bit-reversed [w^i] for i < n/2 (canonical ints)."""
    p = MOD[field]
    w = root_of_unity(field, n.bit_length() - 1)
    if inverse:
        w = pow(w, n - 1, p)
    series = [pow(w, i, p) for i in range(n // 2)]
    bits = (n // 2).bit_length() - 1
    return [series[bitrev(i, bits)] for i in range(n // 2)]


def eval_many(field, coeffs, xs):
    """math/src/polynom: Horner evaluation at every x."""
    p = MOD[field]
    out = []
    for x in xs:
        acc = 0
        for c in reversed(coeffs):
            acc = (acc * x + c) % p
        out.append(acc)
    return out


def lde_domain(field, n, blowup, offset=None):
    p = MOD[field]
    offset = GENERATOR[field] if offset is None else offset
    g = root_of_unity(field, (n * blowup).bit_length() - 1)
    return [offset * pow(g, k, p) % p for k in range(n * blowup)]


def interpolate(field, evals):
    """c_j = n^-1 sum_k v_k w^(-jk)   (math/src/fft/serial.rs:64-73 semantics), O(n^2)."""
    p = MOD[field]
    n = len(evals)
    w = root_of_unity(field, n.bit_length() - 1) if n > 1 else 1
    winv = pow(w, -1, p)
    ninv = pow(n, -1, p)
    return [ninv * sum(v * pow(winv, j * k, p) for k, v in enumerate(evals)) % p for j in range(n)]


def elem_bytes(field, v):
    """bytes that enter the hash for one canonical value (F4/a19): f64 8 B LE, f128 16 B LE."""
    return int(v).to_bytes(8 if field == 0 else 16, "little")


def hash_bytes(hash_id, data):
    if hash_id == 1:
        return hashlib.sha3_256(data).digest()
    return _blake3.blake3(data).digest()


def merkle_nodes(hash_id, leaves):
    """crypto/src/merkle/mod.rs:350-374: heap layout, nodes[0] = zeros, root nodes[1]."""
    n = len(leaves)
    nodes = [bytes(32)] * n
    half = n // 2
    for i in range(half):
        nodes[half + i] = hash_bytes(hash_id, leaves[2 * i] + leaves[2 * i + 1])
    for i in range(half - 1, 0, -1):
        nodes[i] = hash_bytes(hash_id, nodes[2 * i] + nodes[2 * i + 1])
    return nodes


def commit(field, hash_id, traces, blowup):
    """Prover::build_trace_commitment semantics (prover/src/lib.rs:615-670) on canonical ints.
    traces: list of traces, each a list of columns (lists of ints). Returns polys, lde rows, leaves, nodes."""
    n = len(traces[0][0])
    dom = lde_domain(field, n, blowup)
    polys = [[interpolate(field, col) for col in tr] for tr in traces]
    ldes = [[eval_many(field, poly, dom) for poly in tp] for tp in polys]  # [trace][col][row]
    leaves = []
    for k in range(n * blowup):
        data = b"".join(elem_bytes(field, col[k]) for tl in ldes for col in tl)
        leaves.append(hash_bytes(hash_id, data))
    return polys, ldes, leaves, merkle_nodes(hash_id, leaves)


def fri_fold(field, evals, folding, offset, alpha):
    """Definition of the degree-respecting projection (what fri/src/folding/mod.rs:115-146 computes): for every i < n/N
    the polynomial of degree < N through the N points (x_j, evals[i + j*n/N]), x_j = offset * g^(i + j*n/N), evaluated
    at alpha. Canonical ints in, canonical ints out. O(n*N^2) Lagrange form, independent of any FFT."""
    p = MOD[field]
    n = len(evals)
    rows = n // folding
    g = root_of_unity(field, n.bit_length() - 1)
    out = []
    for i in range(rows):
        xs = [offset * pow(g, i + j * rows, p) % p for j in range(folding)]
        ys = [evals[i + j * rows] for j in range(folding)]
        acc = 0
        for j in range(folding):
            num, den = 1, 1
            for m in range(folding):
                if m != j:
                    num = num * (alpha - xs[m]) % p
                    den = den * (xs[j] - xs[m]) % p
            acc = (acc + ys[j] * num * pow(den, -1, p)) % p
        out.append(acc)
    return out
