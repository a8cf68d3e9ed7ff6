"""Frozen golden vectors (tests/golden/golden.json, made by tests/golden/make_golden.py) against
  * the C oracle and the Python big-int ground truth   (CPU, `-m "not gpu"`)
  * the CUDA path through the C ABI                    (`-m gpu`)
Inputs are the reference's fixtures (crypto/src/merkle/tests.rs:12-99 LEAVES4/8, prover/src/trace/tests.rs:28-38 fib
columns) and the SURVEY 8(d) seeded generator. Bit-exact: words, digests, nodes, root."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests import fixtures as fx
from tests.golden.make_golden import HASH_NAME, seeded_inputs, sha, words_hex

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "golden.json")) as _f:
    G = json.load(_f)
HASH_ID = {v: k for k, v in HASH_NAME.items()}
F64, F128 = 0, 1


def words(field, hexes):
    if field == F64:
        return np.array([int(h, 16) for h in hexes], dtype=np.uint64)
    return np.array([(int(h, 16) & fx.MASK, int(h, 16) >> 64) for h in hexes], dtype=np.uint64)


def small_traces(case):
    return [np.stack([words(case["field"], col) for col in tr]) for tr in case["trace"]]


def check_small(case, polys, ldes, leaves, nodes, root):
    f = case["field"]
    for t, tr in enumerate(case["trace"]):
        for c in range(len(tr)):
            assert words_hex(f, polys[t][c]) == case["polys"][t][c], (case["name"], "polys", t, c)
        for k, row in enumerate(case["lde_rows"][t]):
            assert words_hex(f, ldes[t][k, : len(tr)]) == row, (case["name"], "lde row", t, k)
            pad = np.asarray(ldes[t][k, len(tr):])
            assert not pad.any(), (case["name"], "padding columns must be zero", t, k)
    assert [bytes(x).hex() for x in leaves] == case["leaves"], case["name"]
    assert bytes(nodes[1]).hex() == case["root"] and bytes(root).hex() == case["root"], case["name"]


def check_seeded(case, polys, ldes, leaves, nodes, root):
    assert [sha(p) for p in polys] == case["polys_sha256"], case["name"]
    assert [sha(l) for l in ldes] == case["lde_sha256"], case["name"]
    assert sha(leaves) == case["leaves_sha256"], case["name"]
    assert sha(np.asarray(nodes)[1:]) == case["nodes_sha256"], case["name"]
    assert bytes(root).hex() == case["root"], case["name"]


# ------------------------------------------------------------------------------------------------ CPU: the checkers
def test_golden_file_is_self_consistent():
    """root == nested merge of the listed nodes with an implementation that is neither oracle.c nor the kernels."""
    from oracle import pyref

    for m in G["merkle"]:
        h = HASH_ID[m["hash"]]
        leaves = [bytes(x) for x in getattr(fx, m["leaves"])]
        nodes = pyref.merkle_nodes(h, leaves)
        assert [n.hex() for n in nodes] == m["nodes"] and nodes[1].hex() == m["root"]
    # crypto/src/merkle/tests.rs:68-99: LEAVES4 root = merge(merge(l0,l1), merge(l2,l3))
    l = [bytes(x) for x in fx.LEAVES4]
    s3 = lambda b: hashlib.sha3_256(b).digest()
    want = s3(s3(l[0] + l[1]) + s3(l[2] + l[3])).hex()
    assert [m["root"] for m in G["merkle"] if m["leaves"] == "LEAVES4" and m["hash"] == "sha3_256"] == [want]


@pytest.mark.parametrize("m", G["merkle"], ids=lambda m: m["leaves"] + "_" + m["hash"])
def test_oracle_merkle_golden(oracle, m):
    nodes = oracle.merkle_nodes(HASH_ID[m["hash"]], np.array(getattr(fx, m["leaves"]), dtype=np.uint8))
    assert [bytes(r).hex() for r in nodes] == m["nodes"]


@pytest.mark.parametrize("case", G["small"], ids=lambda c: c["name"])
def test_oracle_small_golden(oracle, case):
    got = oracle.commit_trace(case["field"], HASH_ID[case["hash"]], small_traces(case), case["blowup"])
    check_small(case, got["polys"], got["lde"], got["leaves"], got["nodes"], got["root"])


def test_pyref_small_golden():
    """the big-int O(n^2) definition (math/src/fft/tests.rs:19-60 style) reproduces the frozen fib commitment."""
    from oracle import pyref

    case = G["small"][0]
    tr = [[[int(h, 16) for h in col] for col in t] for t in case["trace"]]
    polys, ldes, leaves, nodes = pyref.commit(case["field"], HASH_ID[case["hash"]], tr, case["blowup"])
    assert [[format(v, "032x") for v in col] for col in polys[0]] == case["polys"][0]
    assert [x.hex() for x in leaves] == case["leaves"] and nodes[1].hex() == case["root"]


@pytest.mark.parametrize("case", G["seeded"], ids=lambda c: c["name"])
def test_oracle_seeded_golden(oracle, case):
    got = oracle.commit_trace(case["field"], HASH_ID[case["hash"]], seeded_inputs(case["generator"]), case["blowup"])
    check_seeded(case, got["polys"], got["lde"], got["leaves"], got["nodes"], got["root"])


# ------------------------------------------------------------------------------------------------ GPU: the product
@pytest.fixture(scope="module")
def wg():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


def gpu_commit(wg, field, hash_id, traces, blowup):
    cms = [wg.matrix.ColMatrix(field, list(t)) for t in traces]
    n = cms[0].num_rows()
    ldes, tree, polys = wg.prover.build_trace_commitment(hash_id, cms, wg.matrix.StarkDomain(field, n, blowup))
    return [np.stack(p.columns) for p in polys], [l.data for l in ldes], tree.leaves(), tree.nodes, tree.root()


@pytest.mark.gpu
@pytest.mark.parametrize("m", G["merkle"], ids=lambda m: m["leaves"] + "_" + m["hash"])
def test_gpu_merkle_golden(wg, m):
    tree = wg.merkle.MerkleTree.new(HASH_ID[m["hash"]], np.array(getattr(fx, m["leaves"]), dtype=np.uint8))
    assert [bytes(r).hex() for r in tree.nodes[1:]] == m["nodes"][1:]
    assert tree.root().hex() == m["root"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", G["small"], ids=lambda c: c["name"])
def test_gpu_small_golden(wg, case):
    check_small(case, *gpu_commit(wg, case["field"], HASH_ID[case["hash"]], small_traces(case), case["blowup"]))


@pytest.mark.gpu
@pytest.mark.parametrize("case", G["seeded"], ids=lambda c: c["name"])
def test_gpu_seeded_golden(wg, case):
    check_seeded(case, *gpu_commit(wg, case["field"], HASH_ID[case["hash"]], seeded_inputs(case["generator"]), case["blowup"]))
