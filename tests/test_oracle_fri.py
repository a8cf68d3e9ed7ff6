"""CPU pinning of the FRI-layer part of the oracle (oracle.c or_fri_*) and of the host-side FRI logic.
The reference's FRI tests (fri/src/prover/tests.rs:20-130) only prove-and-verify; they store no digest or folded value.
So the restatement of folding::apply_drp is pinned against the DEFINITION of the degree-respecting projection
(oracle/pyref.py fri_fold: Lagrange interpolation over the coset, no FFT) and against the polynomial identity it
exists for: folding evaluations of f over offset*<g> gives evaluations of sum_k alpha^k f_k over offset^N*<g^N>."""
import numpy as np
import pytest

from oracle import pyref
from tests import fixtures as fx

F64, F128, B3, SHA3 = 0, 1, 0, 1


def rand_ints(field, seed, n):
    p = pyref.MOD[field]
    rng = np.random.default_rng(seed)
    return [int.from_bytes(rng.bytes(16), "little") % p for _ in range(n)]


def one_elem(field, v):
    return pyref.ints_to_array(field, [v])[0] if field == F128 else pyref.ints_to_array(field, [v])


@pytest.mark.parametrize("field", [F64, F128])
@pytest.mark.parametrize("folding", [2, 4, 8, 16])
def test_apply_drp_is_the_coset_interpolant(oracle, field, folding):
    for rows in (2, 8, 32):
        n = rows * folding
        ev = rand_ints(field, 7 * n + folding, n)
        alpha = rand_ints(field, n + 1, 1)[0]
        offset = pyref.GENERATOR[field]
        tr = oracle.fri_commit_layer(field, B3, pyref.ints_to_array(field, ev), folding)["transposed"]
        got = oracle.fri_apply_drp(field, tr, folding, one_elem(field, offset), one_elem(field, alpha))
        assert pyref.array_to_ints(field, got) == pyref.fri_fold(field, ev, folding, offset, alpha), (rows, folding)


@pytest.mark.parametrize("field", [F64, F128])
@pytest.mark.parametrize("folding", [2, 4])
def test_apply_drp_respects_degree(oracle, field, folding):
    """fri/src/prover/tests.rs:58-69 builds evaluations of a low-degree polynomial; after one projection the result must
    be the evaluations of g(y) = sum_k alpha^k f_k(y), f(x) = sum_k x^k f_k(x^N), over the domain offset^N * <w^N>."""
    p = pyref.MOD[field]
    d, blowup = 64, 8
    n = d * blowup
    coeffs = rand_ints(field, 99 + folding, d)
    offset = pyref.GENERATOR[field]
    alpha = rand_ints(field, 5, 1)[0]
    dom = pyref.lde_domain(field, d, blowup, offset)
    ev = pyref.eval_many(field, coeffs, dom)
    tr = oracle.fri_commit_layer(field, B3, pyref.ints_to_array(field, ev), folding)["transposed"]
    got = pyref.array_to_ints(field, oracle.fri_apply_drp(field, tr, folding, one_elem(field, offset), one_elem(field, alpha)))
    folded = [sum(pow(alpha, k, p) * coeffs[t * folding + k] for k in range(folding)) % p for t in range(d // folding)]
    g = pyref.root_of_unity(field, n.bit_length() - 1)
    dom2 = [pow(offset, folding, p) * pow(g, folding * i, p) % p for i in range(n // folding)]
    assert got == pyref.eval_many(field, folded, dom2)


@pytest.mark.parametrize("field,hash_id", [(F64, B3), (F128, B3), (F128, SHA3), (F64, SHA3)])
def test_commit_layer_against_definitions(oracle, field, hash_id):
    for folding in (2, 4, 8, 16):
        rows = 16
        n = rows * folding
        ev = rand_ints(field, 1000 + folding, n)
        got = oracle.fri_commit_layer(field, hash_id, pyref.ints_to_array(field, ev), folding)
        # utils/core/src/lib.rs:206-225: result[i][j] = source[i + j * row_count]
        want_rows = [[ev[i + j * rows] for j in range(folding)] for i in range(rows)]
        assert [pyref.array_to_ints(field, got["transposed"][i]) for i in range(rows)] == want_rows
        leaves = [pyref.hash_bytes(hash_id, b"".join(pyref.elem_bytes(field, v) for v in r)) for r in want_rows]
        assert [bytes(x) for x in got["leaves"]] == leaves
        assert [bytes(x) for x in got["nodes"]][1:] == pyref.merkle_nodes(hash_id, leaves)[1:]


def test_options_and_positions():
    from starkpack_winterfell_b200.fri import FriOptions, fold_positions

    # fri/src/prover/tests.rs:20-48: trace 2^12, blowup 8; folding 2 / remainder degree 7; folding 4 / 255
    assert FriOptions(8, 2, 7).num_fri_layers(1 << 15) == 9
    assert FriOptions(8, 4, 255).num_fri_layers(1 << 15) == 2
    assert FriOptions(8, 4, 255).num_fri_layers(2048) == 0
    # fri/src/folding/mod.rs:160-176: modulo the target domain, first occurrence kept, order preserved
    assert fold_positions([5, 13, 1, 9, 21], 32, 4) == [5, 1]
    assert fold_positions([31, 0, 16], 32, 2) == [15, 0]
    with pytest.raises(AssertionError):
        FriOptions(8, 3, 7)
