"""Fixed inputs taken from the reference's own tests (values only; expected outputs are recomputed by
the reference's tests too, so goldens are frozen by tests/golden/make_golden.py from the pinned oracle)."""
import numpy as np

# crypto/src/merkle/tests.rs:13-66
LEAVES4 = [
    [166, 168, 47, 140, 153, 86, 156, 86, 226, 229, 149, 76, 70, 132, 209, 109, 166, 193, 113, 197, 42, 116, 170, 144, 74, 104, 29, 110, 220, 49, 224, 123],
    [243, 57, 40, 140, 185, 79, 188, 229, 232, 117, 143, 118, 235, 229, 73, 251, 163, 246, 151, 170, 14, 243, 255, 127, 175, 230, 94, 227, 214, 5, 89, 105],
    [11, 33, 220, 93, 26, 67, 166, 154, 93, 7, 115, 130, 70, 13, 166, 45, 120, 233, 175, 86, 144, 110, 253, 250, 67, 108, 214, 115, 24, 132, 45, 234],
    [47, 173, 224, 232, 30, 46, 197, 186, 215, 15, 134, 211, 73, 14, 34, 216, 6, 11, 217, 150, 90, 242, 8, 31, 73, 85, 150, 254, 229, 244, 23, 231],
]
LEAVES8 = [
    [115, 29, 176, 48, 97, 18, 34, 142, 51, 18, 164, 235, 236, 96, 113, 132, 189, 26, 70, 93, 101, 143, 142, 52, 252, 33, 80, 157, 194, 52, 209, 129],
    [52, 46, 37, 214, 24, 248, 121, 199, 229, 25, 171, 67, 65, 37, 98, 142, 182, 72, 202, 42, 223, 160, 136, 60, 38, 255, 222, 82, 26, 27, 130, 203],
    [130, 43, 231, 0, 59, 228, 152, 140, 18, 33, 87, 27, 49, 190, 44, 82, 188, 155, 163, 108, 166, 198, 106, 143, 83, 167, 201, 152, 106, 176, 242, 119],
    [207, 158, 56, 143, 28, 146, 238, 47, 169, 32, 166, 97, 163, 238, 171, 243, 33, 209, 120, 219, 17, 182, 96, 136, 13, 90, 6, 27, 247, 242, 49, 111],
    [179, 64, 123, 119, 226, 139, 161, 127, 36, 251, 218, 88, 20, 217, 212, 85, 112, 85, 185, 193, 230, 181, 4, 22, 54, 219, 135, 98, 235, 180, 182, 7],
    [101, 240, 19, 44, 43, 213, 31, 138, 39, 26, 82, 147, 255, 96, 234, 51, 105, 6, 233, 144, 255, 187, 242, 3, 157, 246, 55, 175, 98, 121, 92, 175],
    [25, 96, 149, 179, 94, 8, 170, 214, 169, 135, 12, 212, 224, 157, 182, 127, 233, 93, 151, 214, 36, 183, 156, 212, 233, 152, 125, 244, 146, 161, 75, 128],
    [247, 43, 130, 141, 234, 172, 61, 187, 109, 31, 56, 30, 14, 232, 92, 158, 48, 161, 108, 234, 170, 180, 233, 77, 200, 248, 45, 152, 125, 11, 1, 171],
]
# prover/src/trace/tests.rs:28-38 (build_fib_trace(16): 8 rows x 2 columns, f128)
FIB_COLS = [[1, 2, 5, 13, 34, 89, 233, 610], [1, 3, 8, 21, 55, 144, 377, 987]]

P64 = 2**64 - 2**32 + 1
P128 = 2**128 - 45 * 2**40 + 1
MASK = 2**64 - 1


def splitmix64(x):
    """vectorised splitmix64 finaliser over uint64 arrays (SURVEY 8(d) synthetic generator)."""
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def synth_f64(seed, cols, n):
    """[cols, n] uniform canonical values < p (as uint64). Callers convert to Montgomery words."""
    with np.errstate(over="ignore"):
        c = np.arange(cols, dtype=np.uint64)[:, None] << np.uint64(32)
        r = np.arange(n, dtype=np.uint64)[None, :]
        x = splitmix64(np.uint64(seed) + c + r)
    return np.where(x >= np.uint64(P64), x - np.uint64(P64), x)


def synth_f128(seed, cols, n):
    """[cols, n, 2] canonical (lo, hi) values < q."""
    with np.errstate(over="ignore"):
        c = np.arange(cols, dtype=np.uint64)[:, None] << np.uint64(32)
        r = np.arange(n, dtype=np.uint64)[None, :]
        lo = splitmix64(np.uint64(seed) + c + r)
        hi = splitmix64(np.uint64(seed) + c + r + np.uint64(1 << 31))
    # reduce: values >= q only when hi == 2^64-1 and lo >= low word of q; clear the top bit instead
    hi = np.where(hi == np.uint64(MASK), hi >> np.uint64(1), hi)
    return np.stack([lo, hi], axis=-1)


def fib_trace_f128(n, start=(1, 1)):
    """prover/src/tests/mod.rs:17-29 recurrence: reg1' = reg1 + reg2; reg2' = reg1' + reg2 (mod q). [2, n, 2]."""
    a, b = start
    out = np.empty((2, n, 2), dtype=np.uint64)
    for i in range(n):
        out[0, i] = (a & MASK, a >> 64)
        out[1, i] = (b & MASK, b >> 64)
        a = (a + b) % P128
        b = (a + b) % P128
    return out
