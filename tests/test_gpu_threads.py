"""SURVEY 8(b) threading: the reference calls the fine-grained functions concurrently from rayon workers
(ColMatrix::interpolate_columns -> par_iter over columns, prover/src/matrix/col_matrix.rs:198-204; segment cosets,
segments.rs:127-137), so the C ABI has to be re-entrant. Eight host threads hammer different entry points at once
(ctypes drops the GIL during the call); every result must equal the oracle's."""
import threading

import numpy as np
import pytest

from tests import fixtures as fx

pytestmark = pytest.mark.gpu
F64, F128, B3, SHA3 = 0, 1, 0, 1


def test_concurrent_entry_points(oracle):
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    gen = w.matrix.GENERATOR_INTERNAL
    jobs = []
    for k in range(8):
        field = F64 if k % 2 == 0 else F128
        n = 1 << (9 + k)  # 2^9 .. 2^16: single-pass and two-pass transforms, different table keys per thread
        cols = fx.synth_f64(40 + k, 3, n) if field == F64 else fx.synth_f128(40 + k, 3, n)
        jobs.append((k, field, n, cols))
    want = {}
    for k, field, n, cols in jobs:
        want[k] = dict(
            poly=oracle.interpolate_poly(field, cols[0]),
            lde=oracle.evaluate_poly_with_offset(field, cols[1], gen[field], 4),
            commit=oracle.commit_trace(field, SHA3 if k % 3 == 0 else B3, [cols], 8, want=("nodes",)),
            fri=oracle.fri_commit_layer(field, B3, cols[2], 4)["root"],
        )
    got, errors = {}, []
    barrier = threading.Barrier(len(jobs))

    def run(k, field, n, cols):
        try:
            barrier.wait()
            out = {}
            for rep in range(3):
                v = cols[0].copy()
                w.fft.interpolate_poly(field, v)
                out["poly"] = v
                out["lde"] = w.fft.evaluate_poly_with_offset(field, cols[1], None, gen[field], 4)
                cm = w.matrix.ColMatrix(field, list(cols))
                _, tree, _ = w.prover.build_trace_commitment(SHA3 if k % 3 == 0 else B3, [cm], w.matrix.StarkDomain(field, n, 8), want_lde=False)
                out["root"] = tree.root()
                out["fri"] = w.fri.commit_layer(field, B3, cols[2], 4)[1].root()
            got[k] = out
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    ts = [threading.Thread(target=run, args=j) for j in jobs]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    for k, field, n, cols in jobs:
        assert np.array_equal(got[k]["poly"], want[k]["poly"]), k
        assert np.array_equal(got[k]["lde"], want[k]["lde"]), k
        assert got[k]["root"] == bytes(want[k]["commit"]["root"]), k
        assert got[k]["fri"] == want[k]["fri"], k
