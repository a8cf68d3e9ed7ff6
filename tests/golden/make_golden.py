"""Freezes golden vectors for the trace-commitment path into tests/golden/golden.json.

Run from the repo root:  python tests/golden/make_golden.py

The reference stores no digest or LDE value in its tests (expected values are recomputed by the same
code), so the goldens are produced here from two independent sources and written only when they agree:
  * "small" cases: oracle/pyref.py (Python big ints, O(n^2) eval_many, the `blake3` module, hashlib.sha3_256) AND
    oracle/oracle.c (C restatement of the reference's algorithm) must give identical words and digests;
  * "seeded" cases (2^10-row cfg-shaped instances): oracle.c only; LDE / leaves / nodes are stored as SHA-256
    checksums of their little-endian bytes, the root in full.
Inputs are the reference's fixtures (tests/fixtures.py: LEAVES4/8, the 8-row fib trace) and the SURVEY 8(d) seeded
generator, so nothing but this script and fixtures.py is needed to regenerate the file.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import pyref  # noqa: E402
from tests import fixtures as fx  # noqa: E402

F64, F128, B3, SHA3 = 0, 1, 0, 1
HASH_NAME = {B3: "blake3_256", SHA3: "sha3_256"}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def words_hex(field, arr):
    """stored words as hex strings (f64: Montgomery word; f128: canonical value)."""
    arr = np.asarray(arr, dtype=np.uint64)
    if field == F64:
        return [format(int(x), "016x") for x in arr.reshape(-1)]
    a = arr.reshape(-1, 2)
    return [format(int(a[i, 0]) | (int(a[i, 1]) << 64), "032x") for i in range(a.shape[0])]


def merkle_cases():
    out = []
    for name, leaves in (("LEAVES4", fx.LEAVES4), ("LEAVES8", fx.LEAVES8)):
        lb = [bytes(x) for x in leaves]
        for h in (B3, SHA3):
            nodes_py = pyref.merkle_nodes(h, lb)
            nodes_c = oracle.merkle_nodes(h, np.array(leaves, dtype=np.uint8))
            assert [bytes(r) for r in nodes_c] == nodes_py, (name, h)
            out.append(dict(leaves=name, hash=HASH_NAME[h], nodes=[n.hex() for n in nodes_py], root=nodes_py[1].hex()))
    return out


def small_commit(name, field, hash_id, traces_int, blowup):
    """traces_int: [trace][col][row] canonical ints."""
    polys, ldes, leaves, nodes = pyref.commit(field, hash_id, traces_int, blowup)
    arrs = [np.stack([pyref.ints_to_array(field, col) for col in tr]) for tr in traces_int]
    got = oracle.commit_trace(field, hash_id, arrs, blowup)
    for t, tr in enumerate(traces_int):
        for c in range(len(tr)):
            assert pyref.array_to_ints(field, got["polys"][t][c]) == polys[t][c], (name, "polys", t, c)
            assert pyref.array_to_ints(field, got["lde"][t][:, c]) == ldes[t][c], (name, "lde", t, c)
    assert [bytes(r) for r in got["leaves"]] == leaves, (name, "leaves")
    assert [bytes(r) for r in got["nodes"]] == nodes, (name, "nodes")
    return dict(
        name=name, field=field, hash=HASH_NAME[hash_id], blowup=blowup,
        trace=[[words_hex(field, a[c]) for c in range(a.shape[0])] for a in arrs],
        polys=[[words_hex(field, got["polys"][t][c]) for c in range(len(tr))] for t, tr in enumerate(traces_int)],
        lde_rows=[[words_hex(field, got["lde"][t][k, : len(tr)]) for k in range(len(tr[0]) * blowup)] for t, tr in enumerate(traces_int)],
        leaves=[x.hex() for x in leaves], root=nodes[1].hex(),
    )


def seeded_commit(name, field, hash_id, traces, blowup, gen):
    got = oracle.commit_trace(field, hash_id, traces, blowup)
    return dict(
        name=name, field=field, hash=HASH_NAME[hash_id], blowup=blowup, generator=gen,
        rows=int(traces[0].shape[1]), cols=[int(t.shape[0]) for t in traces],
        polys_sha256=[sha(p) for p in got["polys"]], lde_sha256=[sha(l) for l in got["lde"]],
        leaves_sha256=sha(got["leaves"]), nodes_sha256=sha(got["nodes"][1:]), root=bytes(got["root"]).hex(),
    )


def mont(canon):
    return np.array([(int(x) << 64) % fx.P64 for x in canon.reshape(-1)], dtype=np.uint64).reshape(canon.shape)


def seeded_inputs(gen):
    """the generator strings stored in golden.json; tests/test_golden.py evaluates the same table."""
    kind, *a = gen.split(":")
    if kind == "synth_f64":  # seed, cols, rows; canonical draw stored as Montgomery word
        return [mont(fx.synth_f64(int(a[0], 16), int(a[1]), int(a[2])))]
    if kind == "synth_f128":
        return [fx.synth_f128(int(a[0], 16), int(a[1]), int(a[2]))]
    if kind == "fib_f128":  # rows, n_traces; trace t starts from (t+1, t+1)
        return [fx.fib_trace_f128(int(a[0]), (t + 1, t + 1)) for t in range(int(a[1]))]
    raise ValueError(gen)


def main():
    oracle.build()
    g = dict(about="frozen by tests/golden/make_golden.py; see its docstring", merkle=merkle_cases(), small=[], seeded=[])
    fib = [[list(c) for c in fx.FIB_COLS]]
    g["small"].append(small_commit("fib_8x2_f128_b8_blake3", F128, B3, fib, 8))
    g["small"].append(small_commit("fib_8x2_f128_b8_sha3", F128, SHA3, fib, 8))
    s64 = fx.synth_f64(0xC0FFEE02, 3, 16)
    g["small"].append(small_commit("synth_16x3_f64_b8_blake3", F64, B3, [[[int(x) for x in c] for c in s64]], 8))
    g["small"].append(small_commit("synth_16x3_f64_b4_sha3", F64, SHA3, [[[int(x) for x in c] for c in s64]], 4))
    two = [[[int(x) for x in c] for c in fx.synth_f64(0xC0FFEE10 + t, 2, 8)] for t in range(2)]
    g["small"].append(small_commit("agg2_8x2_f64_b8_blake3", F64, B3, two, 8))
    for name, field, h, blowup, gen in [
        ("cfg2_shape_2^10x100_f64_blake3", F64, B3, 8, "synth_f64:C0FFEE02:100:1024"),
        ("cfg5_shape_2^10x256_f64_blake3", F64, B3, 8, "synth_f64:C0FFEE05:256:1024"),
        ("cfg2_shape_2^13x10_f64_sha3", F64, SHA3, 8, "synth_f64:C0FFEE02:10:8192"),
        ("cfg1_shape_2^10x2_f128_blake3", F128, B3, 8, "fib_f128:1024:1"),
        ("cfg3_shape_2^10x4_f128_sha3", F128, SHA3, 8, "synth_f128:C0FFEE03:4:1024"),
        ("cfg4_shape_8x2^10x2_f128_blake3", F128, B3, 8, "fib_f128:1024:8"),
    ]:
        g["seeded"].append(seeded_commit(name, field, h, seeded_inputs(gen), blowup, gen))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")
    with open(path, "w") as f:
        json.dump(g, f, indent=0, sort_keys=True)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
