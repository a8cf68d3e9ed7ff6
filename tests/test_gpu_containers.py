"""The containers Prover::build_trace_commitment fills and the calls later prover steps make on them
(SURVEY 8(a) row a18, 8(b)): TraceLde frame reads (prover/src/trace/trace_lde.rs:78-108) from a host RowMatrix and from
a resident commitment, TracePolyTable::evaluate_at / get_ood_frame (prover/src/trace/poly_table.rs:66-80 ->
ColMatrix::evaluate_columns_at, col_matrix.rs:210-216), RowMatrix::commit_to_comb_rows over already-built LDE matrices
(prover/src/matrix/row_matrix.rs:204-238)."""
import numpy as np
import pytest

from oracle import pyref
from tests import fixtures as fx

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wg():
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    return w


@pytest.mark.parametrize("field", [0, 1])
def test_trace_lde_frames_host_and_resident(wg, oracle, field):
    from starkpack_winterfell_b200.prover import EvaluationFrame, ResidentCommitment, TraceLde

    n, blowup, cols = 256, 8, 5
    tr = fx.synth_f64(3, cols, n) if field == 0 else fx.synth_f128(3, cols, n)
    want = oracle.commit_trace(field, 0, [tr], blowup)
    cm = wg.matrix.ColMatrix(field, list(tr))
    dom = wg.matrix.StarkDomain(field, n, blowup)
    ldes, tree, polys = wg.prover.build_trace_commitment(0, [cm], dom)
    rc = ResidentCommitment.from_traces(0, [cm], dom)
    for lde in (TraceLde(ldes[0], blowup), TraceLde(rc, blowup)):
        assert (lde.trace_len(), lde.blowup(), lde.main_trace_width(), lde.aux_trace_width()) == (n * blowup, blowup, cols, 0)
        frame = EvaluationFrame(cols, field)
        for step in (0, 1, n * blowup - blowup, n * blowup - 1):  # the last ones wrap around (trace_lde.rs:83-84)
            lde.read_main_trace_frame_into(step, frame)
            assert np.array_equal(frame.current, want["lde"][0][step][:cols])
            assert np.array_equal(frame.next, want["lde"][0][(step + blowup) % (n * blowup)][:cols])
    tc = wg.prover.TraceCommitment(ldes, tree, [blowup])
    assert tc.root() == bytes(want["root"])
    rc.close()


@pytest.mark.parametrize("field", [0, 1])
def test_trace_poly_table_evaluate_at(wg, oracle, field):
    from starkpack_winterfell_b200.prover import TracePolyTable

    p = pyref.MOD[field]
    for n, cols in ((8, 2), (256, 3), (1000 // 8 * 8 + 24, 1), (4096, 4)):
        n = 1 << (n.bit_length() - 1)
        tr = fx.synth_f64(n, cols, n) if field == 0 else fx.synth_f128(n, cols, n)
        polys = wg.matrix.ColMatrix(field, list(tr)).interpolate_columns()
        table = TracePolyTable(polys)
        assert table.poly_size() == n and table.num_main_trace_polys() == cols
        z = 0x123456789ABCDEF % p
        g = pyref.root_of_unity(field, n.bit_length() - 1)
        zi, zgi = pyref.ints_to_array(field, [z])[0], pyref.ints_to_array(field, [z * g % p])[0]
        frame = table.get_ood_frame(zi, zgi)
        for got, x in zip(frame, (z, z * g % p)):
            coeffs = [pyref.array_to_ints(field, c) for c in polys.columns]
            want = [sum(c * pow(x, i, p) for i, c in enumerate(cs)) % p for cs in coeffs]
            assert pyref.array_to_ints(field, got) == want
        assert np.array_equal(table.get_main_trace_poly(0), polys.columns[0])


@pytest.mark.parametrize("field,hash_id", [(0, 0), (1, 1)])
def test_commit_to_comb_rows_over_built_ldes(wg, oracle, field, hash_id):
    """T separately built LDE matrices of different widths -> one tree (row_matrix.rs:204-238)."""
    n, blowup = 512, 4
    widths = [3, 8, 13, 1]
    trs = [fx.synth_f64(40 + i, c, n) if field == 0 else fx.synth_f128(40 + i, c, n) for i, c in enumerate(widths)]
    want = oracle.commit_trace(field, hash_id, trs, blowup, want=("leaves", "nodes"))
    ldes = [wg.matrix.RowMatrix.evaluate_polys(wg.matrix.ColMatrix(field, list(t)).interpolate_columns(), blowup) for t in trs]
    tree = ldes[0].commit_to_comb_rows(hash_id, ldes)
    assert np.array_equal(tree.leaves(), want["leaves"]) and np.array_equal(tree.nodes, want["nodes"])
    assert tree.root() == bytes(want["root"])
    with pytest.raises(AssertionError):
        ldes[1].commit_to_comb_rows(hash_id, ldes)  # self must be traces_lde[0]


@pytest.mark.parametrize("field,shapes,n", [(1, [2] * 8, 4096), (0, [3, 1, 5, 2], 8192), (0, [2, 2], 256)])
def test_narrow_traces_extended_together(wg, oracle, field, shapes, n):
    """StarkPack commits several narrow traces at once (examples/src/do_work: 8 traces of 2 columns). With the scratch of
    wg_commit_scratch_elems the device entry extends all columns together (full column tiles) and lays every trace out as its
    own matrix afterwards; with the per-trace scratch it extends them one by one. Same polys, LDE matrices (pad columns
    zero), leaves and tree as the oracle either way."""
    import torch

    from starkpack_winterfell_b200._lib import lib
    from starkpack_winterfell_b200.device import TraceCommitter

    blowup = 8
    traces = [(fx.synth_f64 if field == 0 else fx.synth_f128)(700 + 3 * i + c, c, n) for i, c in enumerate(shapes)]
    want = oracle.commit_trace(field, 0, traces, blowup)
    allc = np.concatenate([np.ascontiguousarray(t) for t in traces], axis=0).view(np.int64)
    launches = []
    for per_trace_scratch in (None, max(shapes)):
        tc = TraceCommitter(field, 0, n, blowup, shapes, device="cuda:0", lde_scratch_cols=per_trace_scratch)
        tc.cols.copy_(torch.from_numpy(allc if field == 1 else allc[..., None]))
        l0 = lib().wg_launch_count()
        tc.commit()
        torch.cuda.synchronize()
        launches.append(lib().wg_launch_count() - l0)
        assert tc.root() == bytes(want["root"])
        assert np.array_equal(tc.nodes.cpu().numpy(), want["nodes"])
        for t in range(len(shapes)):
            got = tc.ldes[t].cpu().numpy().view(np.uint64)
            assert np.array_equal(got.reshape(want["lde"][t].shape), want["lde"][t]), (per_trace_scratch, t)
    assert launches[0] != launches[1]  # the two plans really differ
