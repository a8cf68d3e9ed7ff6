"""world_size-2/4 gloo runs of the multi-GPU commit logic (starkpack_winterfell_b200/parallel.py) on CPU: column
partition, the all-to-all into row blocks (uneven shard widths), subtree + top-level Merkle and heap stitching.
Compute is done by a CHECKER backend built on the oracle (tests may use it; the product backend is the C ABI), and
the result must equal the single-process oracle commit of the whole trace."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import fixtures as fx


class OracleBackend:
    def __init__(self, field, hash_id):
        import oracle

        self.o, self.field, self.hash_id = oracle, field, hash_id
        self.w = 1 if field == 0 else 2

    def empty(self, shape, dtype=torch.int64):
        return torch.empty(shape, dtype=dtype)

    def _np(self, t):
        a = t.numpy().view(np.uint64)
        return a[..., 0] if self.field == 0 else a

    def _t(self, a):
        a = np.ascontiguousarray(a).view(np.int64)
        return torch.from_numpy(a[..., None] if self.field == 0 else a)

    def interpolate_lde(self, cols, n, blowup):
        r = self.o.commit_trace(self.field, self.hash_id, [self._np(cols)], blowup, want=("polys", "lde"))
        return self._t(r["polys"][0]), self._t(r["lde"][0])

    def interpolate(self, cols, n):
        return self._t(np.stack([self.o.interpolate_poly(self.field, c) for c in self._np(cols)]))

    def lde(self, polys, n, blowup, out=None, scratch=None):
        p = self._np(polys)
        off = np.array([(7 << 64) % fx.P64], dtype=np.uint64) if self.field == 0 else np.array([3, 0], dtype=np.uint64)
        cols = [self.o.evaluate_poly_with_offset(self.field, c, off, blowup) for c in p]
        rw = (len(cols) + 7) // 8 * 8
        m = np.zeros((n * blowup, rw) + ((2,) if self.field == 1 else ()), dtype=np.uint64)
        for i, c in enumerate(cols):
            m[:, i] = c
        return self._t(m)

    def hash_rows(self, mats, elems, rows):
        return torch.from_numpy(self.o.hash_comb_rows(self.field, self.hash_id, [self._np(m) for m in mats], elems))

    def merkle(self, leaves):
        return torch.from_numpy(self.o.merkle_nodes(self.hash_id, leaves.numpy()))


def _worker(rank, world, port, field, hash_id, n, blowup, total_cols, seed, q, pipeline=0):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from starkpack_winterfell_b200.parallel import DistributedCommitter

        full = fx.synth_f64(seed, total_cols, n) if field == 0 else fx.synth_f128(seed, total_cols, n)
        be = OracleBackend(field, hash_id)
        dc = (DistributedCommitter.auto(be, n, blowup, total_cols) if pipeline < 0
              else DistributedCommitter(be, n, blowup, total_cols, pipeline=pipeline))
        if pipeline < 0:
            assert not dc.fused and dc.pipeline == (4 if total_cols // world >= 64 else 0)
        local = be._t(full[dc.col_start: dc.col_start + dc.col_count])
        root = dc.commit(local)
        nodes = dc.gather_nodes().numpy()
        leaves = dc.gather_leaves().numpy()
        if rank == 0:
            q.put((root, nodes, leaves))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,field,hash_id,n,blowup,total_cols,pipeline", [
    (2, 0, 0, 64, 8, 10, 0),    # even split 5/5
    (2, 0, 0, 32, 4, 21, 0),    # uneven widths: 11 -> rw 16, 10 -> rw 16
    (2, 0, 1, 32, 2, 9, 0),     # 5 -> rw 8, 4 -> rw 8, SHA3
    (4, 1, 0, 16, 8, 6, 0),     # f128, shards of 2,2,1,1 columns
    (2, 0, 0, 16, 8, 17, 0),    # 9 -> rw 16, 8 -> rw 8 (different padded widths across ranks)
    (2, 0, 0, 16, 8, 37, 2),    # pipelined exchange: chunks 16+3 / 16+2
    (2, 0, 0, 16, 4, 21, 4),    # 11 -> chunks 8+3+0+0, 10 -> 8+2+0+0 (empty chunks skipped on every rank)
    (4, 1, 0, 16, 8, 6, 2),     # f128, single-column shards: second chunk empty everywhere
    (2, 0, 0, 16, 2, 130, -1),  # DistributedCommitter.auto: 65 columns per rank -> 4 chunks; no peer memory on CPU
    (2, 0, 0, 16, 8, 10, -1),   # auto on a narrow shard without peer memory: plain all-to-all
])
def test_distributed_commit_matches_single_process(oracle, world, field, hash_id, n, blowup, total_cols, pipeline):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, field, hash_id, n, blowup, total_cols, 77, q, pipeline)) for r in range(world)]
    for p in procs:
        p.start()
    root, nodes, leaves = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = fx.synth_f64(77, total_cols, n) if field == 0 else fx.synth_f128(77, total_cols, n)
    want = oracle.commit_trace(field, hash_id, [full], blowup)
    assert np.array_equal(leaves, want["leaves"])
    assert np.array_equal(nodes, want["nodes"])
    assert root == bytes(want["root"])


def test_split_and_stitch_helpers():
    from starkpack_winterfell_b200.parallel import split_columns

    assert split_columns(100, 8) == [(0, 13), (13, 13), (26, 13), (39, 13), (52, 12), (64, 12), (76, 12), (88, 12)]
    assert split_columns(256, 4) == [(0, 64), (64, 64), (128, 64), (192, 64)]
    from starkpack_winterfell_b200.parallel import split_chunks

    assert split_chunks(128, 4) == [(0, 32), (32, 32), (64, 32), (96, 32)]
    assert split_chunks(13, 4) == [(0, 8), (8, 5), (13, 0), (13, 0)]
    assert split_chunks(50, 2) == [(0, 32), (32, 18)]
