"""The sharded commit behind the C ABI (wg_sharded_*, include/winter_gpu.h) on ONE device: wg_sharded_commit_emulated runs
the per-rank code of all G ranks (LDE per column block, the exchange in every form, row hashing over the received
blocks, subtree, top levels, heap stitching) with device copies standing in for NCCL, and must reproduce the oracle's
single-process commit of the whole trace (Prover::build_trace_commitment, prover/src/lib.rs:615-670; whole-row leaves,
prover/src/matrix/row_matrix.rs:204-238). The real NCCL / P2P run is tests/run_distributed_gpu.py (>= 2 GPUs)."""
import numpy as np
import pytest
import torch

from tests import fixtures as fx

pytestmark = pytest.mark.gpu


def _to_dev(a, field, dev):
    a = np.ascontiguousarray(a).view(np.int64)
    return torch.from_numpy(a[..., None] if field == 0 else a).to(dev)


CASES = [
    # G, field, hash, n, blowup, total_cols
    (2, 0, 0, 8192, 8, 21), (4, 0, 0, 1024, 8, 100), (8, 1, 0, 4096, 8, 16), (2, 0, 1, 256, 4, 17), (8, 0, 0, 2048, 8, 256),
    (2, 0, 0, 1 << 14, 8, 100), (4, 1, 1, 512, 8, 9), (8, 0, 0, 1 << 13, 2, 8), (1, 0, 0, 4096, 8, 13), (2, 1, 0, 64, 2, 2),
]


@pytest.mark.parametrize("G,field,hash_id,n,blowup,total_cols", CASES)
@pytest.mark.parametrize("exchange,chunks", [(1, 0), (2, 2), (2, 4), (3, 0), (4, 2), (0, 0)])
def test_emulated_sharded_commit_matches_oracle(oracle, G, field, hash_id, n, blowup, total_cols, exchange, chunks):
    from starkpack_winterfell_b200 import _lib, sharded

    _lib.init(0)
    dev = torch.device("cuda:0")
    full = fx.synth_f64(5, total_cols, n) if field == 0 else fx.synth_f128(5, total_cols, n)
    root, leaves, nodes = sharded.commit_emulated(field, hash_id, n, blowup, total_cols, G, _to_dev(full, field, dev), exchange, chunks)
    want = oracle.commit_trace(field, hash_id, [full], blowup, want=("leaves", "nodes"))
    assert np.array_equal(leaves, want["leaves"])
    assert np.array_equal(nodes, want["nodes"])
    assert root == bytes(want["root"])


def test_emulated_aggregated_traces_equal_comb_rows(oracle):
    """cfg4's shape: T traces of 2 columns each, one trace (or two) per rank: the global column list hashed row by row is
    commit_to_comb_rows over the T traces (row_matrix.rs:204-238)."""
    from starkpack_winterfell_b200 import _lib, sharded

    _lib.init(0)
    dev = torch.device("cuda:0")
    n = 2048
    traces = [fx.fib_trace_f128(n, start=(t + 1, t + 1)) for t in range(8)]
    want = oracle.commit_trace(1, 0, traces, 8, want=("nodes",))
    full = np.concatenate(traces, axis=0)
    for G in (2, 4, 8):
        root, _, nodes = sharded.commit_emulated(1, 0, n, 8, 16, G, _to_dev(full, 1, dev), want_leaves=False)
        assert root == bytes(want["root"]) and np.array_equal(nodes, want["nodes"]), G


def test_single_rank_handle_accessors(oracle):
    """wg_sharded_commit_dev with world = 1 (no communicator needed) and the handle's accessors."""
    from starkpack_winterfell_b200 import _lib, sharded

    sharded.dist_init(0, 1, device=0)
    dev = torch.device("cuda:0")
    n, blowup, cols = 4096, 8, 21
    full = fx.synth_f64(11, cols, n)
    want = oracle.commit_trace(0, 0, [full], blowup)
    d = _to_dev(full, 0, dev)
    polys = torch.empty_like(d)
    h = sharded.commit_dev(0, 0, n, blowup, cols, d, polys)
    assert h.root() == bytes(want["root"])
    assert (h.lde_rows, h.row_start, h.row_count) == (n * blowup, 0, n * blowup)
    assert h.block_cols == [cols] and h.block_row_widths == [cols]  # exchanged blocks carry no pad columns
    assert np.array_equal(polys.cpu().numpy().view(np.uint64)[:, :, 0], want["polys"][0])
    pos = [0, 1, 77, n * blowup - 1]
    assert np.array_equal(h.rows(pos), want["lde"][0][pos][:, :cols])
    assert np.array_equal(h.digests(0, pos), want["leaves"][pos])
    assert np.array_equal(h.digests(1, [1, 2, 3, n * blowup - 1]), want["nodes"][[1, 2, 3, n * blowup - 1]])
    lv, nd = h.gather()
    assert np.array_equal(lv, want["leaves"]) and np.array_equal(nd, want["nodes"])
    mats, lv2, nd2 = h.download()
    assert np.array_equal(mats[0], want["lde"][0][:, :cols]) and np.array_equal(lv2, want["leaves"]) and np.array_equal(nd2, want["nodes"])
    ph, sent = h.phases_ms()
    assert sent == 0 and set(ph) == set(sharded.PHASE_NAMES) and all(v >= 0 for v in ph.values())
    with pytest.raises(_lib.WinterGpuError):
        h.rows([n * blowup])
    h.close()
    # host-pointer variant
    cols_list = [np.ascontiguousarray(c) for c in full]
    polys_list = [np.empty(n, dtype=np.uint64) for _ in range(cols)]
    h2 = sharded.commit_host(0, 0, n, blowup, cols, cols_list, polys_out=polys_list)
    assert h2.root() == bytes(want["root"]) and np.array_equal(np.stack(polys_list), want["polys"][0])
    h2.close()
    sharded.dist_shutdown()


def test_sharded_argument_errors():
    from starkpack_winterfell_b200 import _lib, sharded

    _lib.init(0)
    dev = torch.device("cuda:0")
    d = torch.zeros((4, 64, 1), dtype=torch.int64, device=dev)
    for kw in (dict(world=3), dict(world=8), dict(n=48)):  # not a power of two / more ranks than columns / rows not pow2
        a = dict(world=2, n=64)
        a.update(kw)
        with pytest.raises(_lib.WinterGpuError):
            sharded.commit_emulated(0, 0, a["n"], 8, 4, a["world"], d)
    with pytest.raises(_lib.WinterGpuError):
        sharded.shard_plan(100, 16, _lib.EXCHANGE_FUSED_P2P)


@pytest.mark.parametrize("cfg", ["cfg1", "cfg3", "cfg4"])
def test_full_size_small_configs_sharded_vs_golden(cfg):
    """BASELINE.json configs 1, 3, 4 at FULL size, as 1 rank and as emulated multi-rank commits, against the roots the CPU
    oracle produced at full size (tests/golden/full_roots.json, made by tests/golden/make_full_roots.py)."""
    import json
    import os

    from starkpack_winterfell_b200 import _lib, sharded

    with open(os.path.join(os.path.dirname(__file__), "golden", "full_roots.json")) as f:
        gold = json.load(f)[cfg]
    _lib.init(0)
    dev = torch.device("cuda:0")
    n = gold["rows"]
    if cfg == "cfg1":
        full = fx.fib_trace_f128(n)
    elif cfg == "cfg3":
        full = fx.synth_f128(0xC0FFEE03, 4, n)
    else:
        full = np.concatenate([fx.fib_trace_f128(n, start=(t + 1, t + 1)) for t in range(8)], axis=0)
    d = _to_dev(full, 1, dev)
    tc = full.shape[0]
    for G in (1, 2) if cfg != "cfg4" else (1, 8):
        root, _, _ = sharded.commit_emulated(1, gold["hash"], n, 8, tc, G, d, want_leaves=False, want_nodes=False)
        assert root.hex() == gold["root"], (cfg, G)
