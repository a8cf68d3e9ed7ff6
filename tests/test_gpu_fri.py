"""GPU parity for the FRI-layer entry points against the CPU oracle's restatement of
FriProver::build_layer / folding::apply_drp (fri/src/prover/mod.rs:185-223, fri/src/folding/mod.rs:115-146). Bit-exact.
Shapes follow fri/src/prover/tests.rs:20-69 (trace 2^12, blowup 8, folding 2 and 4) plus every folding factor, both
fields and hashes, and BASELINE config 1's LDE domain (2^23 f128)."""
import hashlib

import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu
F64, F128, B3, SHA3 = 0, 1, 0, 1


@pytest.fixture(scope="module")
def wg():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


def rand_elems(field, seed, n):
    return fx.synth_f64(seed, 1, n)[0] if field == F64 else fx.synth_f128(seed, 1, n)[0]


def one(field, seed):
    return rand_elems(field, seed, 4)[3:4].reshape(-1)


class HashChannel:
    """deterministic stand-in for the Fiat-Shamir channel (out of scope): alpha = reduced digest of the transcript."""

    def __init__(self, field):
        self.field, self.commitments, self.state = field, [], b"fri-test"

    def commit_fri_layer(self, root):
        self.commitments.append(bytes(root))
        self.state = hashlib.sha256(self.state + bytes(root)).digest()

    def draw_fri_alpha(self):
        self.state = hashlib.sha256(self.state + b"alpha").digest()
        v = int.from_bytes(self.state, "little")
        if self.field == F64:
            return np.array([v % fx.P64], dtype=np.uint64)
        v %= fx.P128
        return np.array([v & fx.MASK, v >> 64], dtype=np.uint64)


@pytest.mark.parametrize("field", [F64, F128])
@pytest.mark.parametrize("hash_id", [B3, SHA3])
def test_commit_layer_and_drp_host(wg, oracle, field, hash_id):
    off = wg.matrix.GENERATOR_INTERNAL[field]
    for folding in (2, 4, 8, 16):
        for rows in (2, 4, 64, 1024, 1 << 14):
            n = rows * folding
            ev = rand_elems(field, 31 * n + folding, n)
            want = oracle.fri_commit_layer(field, hash_id, ev, folding)
            tr, tree = wg.fri.commit_layer(field, hash_id, ev, folding)
            assert np.array_equal(tr, want["transposed"]), (folding, rows)
            assert np.array_equal(tree.leaves(), want["leaves"]), (folding, rows)
            assert np.array_equal(tree.nodes[1:], want["nodes"][1:]) and tree.root() == want["root"], (folding, rows)
            alpha = one(field, n)
            for o in (off, one(field, n + 5)):
                got = wg.fri.apply_drp(field, tr, o, alpha)
                assert np.array_equal(got, oracle.fri_apply_drp(field, want["transposed"], folding, o, alpha)), (folding, rows)


def test_error_codes(wg):
    """prover/mod.rs:178-184 unimplemented folding factor; utils/core/src/lib.rs:208-214 length not divisible;
    MerkleTree::new errors (crypto/src/merkle/mod.rs:118-123)."""
    from starkpack_winterfell_b200._lib import ERR_ARG, ERR_NOT_POW2, ERR_TOO_FEW, ERR_UNSUPPORTED, ERR_ZERO_OFFSET

    ev = rand_elems(F64, 1, 48)
    for ff, n, code in ((3, 48, ERR_UNSUPPORTED), (32, 32, ERR_UNSUPPORTED), (4, 46, ERR_ARG), (4, 4, ERR_TOO_FEW), (4, 48, ERR_NOT_POW2)):
        with pytest.raises(wg.WinterGpuError) as e:
            wg.fri.commit_layer(F64, B3, ev[:n], ff)
        assert e.value.code == code, (ff, n)
    tr = ev[:32].reshape(8, 4)
    with pytest.raises(wg.WinterGpuError) as e:
        wg.fri.apply_drp(F64, tr, np.zeros(1, dtype=np.uint64), one(F64, 3))
    assert e.value.code == ERR_ZERO_OFFSET
    with pytest.raises(wg.WinterGpuError) as e:
        wg.fri.apply_drp(F64, tr, one(F64, 2), np.array([fx.P64], dtype=np.uint64))  # not a reduced element
    assert e.value.code == ERR_ARG


def reference_chain(oracle, field, hash_id, options, ev):
    """FriProver::build_layers (prover/mod.rs:171-223) replayed with the oracle and the same channel."""
    ch = HashChannel(field)
    off = options.domain_offset(field)
    N = options.folding_factor
    layers = []
    for _ in range(options.num_fri_layers(ev.shape[0])):
        l = oracle.fri_commit_layer(field, hash_id, ev, N)
        ch.commit_fri_layer(l["root"])
        ev = oracle.fri_apply_drp(field, l["transposed"], N, off, ch.draw_fri_alpha())
        layers.append(l)
    n = ev.shape[0]
    # interpolate_poly_with_offset (math/src/fft/mod.rs:362-400) = interpolate_poly, then coefficient k *= offset^-k
    poly = oracle.interpolate_poly(field, ev)
    p = fx.P64 if field == F64 else fx.P128
    from oracle import pyref

    ints = pyref.array_to_ints(field, poly)
    oi = pow(pyref.GENERATOR[field], -1, p)
    ints = [c * pow(oi, k, p) % p for k, c in enumerate(ints)]
    rem = pyref.ints_to_array(field, ints[: n // options.blowup_factor])
    digest = oracle.hash_rows(field, hash_id, rem, 1, rem.shape[0], rem.shape[0])[0]
    ch.commit_fri_layer(bytes(digest))
    return layers, rem, ch


@pytest.mark.parametrize("field,hash_id,trace_len,folding,rem_deg", [
    (F128, B3, 1 << 12, 2, 7),      # fri/src/prover/tests.rs:20-32
    (F128, B3, 1 << 12, 4, 255),    # :34-48
    (F64, B3, 1 << 12, 8, 31),
    (F64, SHA3, 1 << 10, 16, 7),
    (F128, SHA3, 1 << 11, 4, 15),
])
def test_fri_prover_layers_and_proof(wg, oracle, field, hash_id, trace_len, folding, rem_deg):
    blowup = 8
    opts = wg.fri.FriOptions(blowup, folding, rem_deg)
    # build_evaluations (tests.rs:58-69): a degree < trace_len polynomial evaluated over the LDE domain (no offset there;
    # here over the coset, as the STARK prover hands it over)
    coeffs = rand_elems(field, 77 + folding, trace_len)
    ev = oracle.evaluate_poly_with_offset(field, coeffs, opts.domain_offset(field), blowup)
    want_layers, want_rem, want_ch = reference_chain(oracle, field, hash_id, opts, ev)

    ch = HashChannel(field)
    prover = wg.fri.FriProver(field, hash_id, opts)
    prover.build_layers(ch, ev)
    assert ch.commitments == want_ch.commitments
    assert prover.num_layers() == len(want_layers) == opts.num_fri_layers(ev.shape[0])
    assert np.array_equal(prover.remainder_poly, want_rem)
    for l, wl in zip(prover.layers, want_layers):
        t = l.tree()
        assert np.array_equal(t.nodes[1:], wl["nodes"][1:]) and np.array_equal(t.leaves(), wl["leaves"])

    positions = [int(x) % ev.shape[0] for x in fx.splitmix64(np.arange(40, dtype=np.uint64))]
    positions = list(dict.fromkeys(positions))[:32]
    host_layers = prover.layers
    proof_layers, rem = prover.build_proof(positions)
    assert prover.num_layers() == 0 and np.array_equal(rem, want_rem)
    dom = ev.shape[0]
    pos = positions
    for pl, l, wl in zip(proof_layers, host_layers, want_layers):
        pos = wg.fri.fold_positions(pos, dom, folding)
        tree = wg.merkle.MerkleTree.from_raw_parts(hash_id, wl["nodes"], wl["leaves"])
        assert pl.proof == tree.prove_batch(pos)
        assert np.array_equal(pl.values.reshape(len(pos), -1), wl["transposed"][pos].reshape(len(pos), -1))
        dom //= folding


def test_fri_cfg1_domain(wg, oracle):
    """BASELINE config 1's LDE domain: 2^23 f128 evaluations, folding 4 — first layer and its projection."""
    n = 1 << 23
    ev = rand_elems(F128, 0xF1, n)
    want = oracle.fri_commit_layer(F128, B3, ev, 4)
    tr, tree = wg.fri.commit_layer(F128, B3, ev, 4)
    assert tree.root() == want["root"] and np.array_equal(tree.nodes[1:], want["nodes"][1:])
    alpha, off = one(F128, 9), wg.matrix.GENERATOR_INTERNAL[F128]
    got = wg.fri.apply_drp(F128, tr, off, alpha)
    assert np.array_equal(got, oracle.fri_apply_drp(F128, want["transposed"], 4, off, alpha))
