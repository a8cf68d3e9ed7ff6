"""torchrun entry (>= 2 GPUs): the sharded commit behind the C ABI over real NCCL / NVLink P2P.

default   small shapes, every exchange form (all-to-all, pipelined x2 / x4, fused P2P, auto), both fields and hashes:
          root, stitched leaves and stitched nodes (wg_sharded_gather) equal to the CPU oracle's single-process commit;
          rows / digests accessors; a second commit on reused buffers gives the same tree. Also the torch.distributed
          DistributedCommitter (parallel.py) on the same inputs. Prints DISTRIBUTED_OK.
--cfg5    BASELINE.json configs[4] at full size (2^24 x 256 f64 on >= 4 GPUs, 2^22 rows on 2): no host can hold its 275 GB LDE,
          so size-independent properties are checked with independent tools (Python big-int Horner, the `blake3` module), as
          tests/test_gpu_fullsize.py does for cfg2: polys evaluate back to the trace, sampled LDE entries of columns owned by
          OTHER ranks equal P_c(offset * g^k) after the exchange, sampled leaves equal blake3 of the 2,048 canonical row bytes
          (two chunks + parent), Merkle paths climb through the local subtree, the peers' subtree roots and the top levels to
          the root, every rank reports the same root. Prints one JSON record (kept under profiles/).
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import fixtures as fx  # noqa: E402

P64 = fx.P64
RINV = pow(2**64, -1, P64)


def _to_dev(a, field, dev):
    a = np.ascontiguousarray(a).view(np.int64)
    return torch.from_numpy(a[..., None] if field == 0 else a).to(dev)


def small_shapes(dev):
    import oracle
    from starkpack_winterfell_b200 import _lib, sharded
    from starkpack_winterfell_b200.parallel import DeviceBackend, DistributedCommitter

    rank, world = dist.get_rank(), dist.get_world_size()
    for field, hash_id, n, blowup, total_cols in [(0, 0, 8192, 8, 100), (1, 0, 4096, 8, 16), (0, 1, 1024, 8, 21), (0, 0, 1 << 14, 8, 256)]:
        if total_cols < world:
            continue
        full = fx.synth_f64(9, total_cols, n) if field == 0 else fx.synth_f128(9, total_cols, n)
        want = oracle.commit_trace(field, hash_id, [full], blowup, want=("leaves", "nodes", "lde")) if rank == 0 else None
        s0, cnt = sharded.shard_columns(total_cols, world, rank)
        cols = _to_dev(full[s0: s0 + cnt], field, dev)
        for exchange, chunks in ((1, 0), (2, 2), (2, 4), (3, 0), (4, 1), (4, 3), (0, 0)):
            if exchange == 3 and world > 8:
                continue
            roots = []
            for rep in range(2):  # second commit reuses the parked buffers / peer mappings
                h = sharded.commit_dev(field, hash_id, n, blowup, total_cols, cols, exchange=exchange, chunks=chunks)
                roots.append(h.root())
                lv, nd = h.gather()
                pos = [h.row_start, h.row_start + h.row_count - 1, h.row_start + h.row_count // 3]
                rows = h.rows(pos)
                dg = h.digests(0, [p - h.row_start for p in pos])
                allrows = [None] * world
                dist.all_gather_object(allrows, (pos, rows, dg))
                if rank == 0:
                    assert h.root() == bytes(want["root"]), (field, n, exchange, chunks)
                    assert np.array_equal(lv, want["leaves"]) and np.array_equal(nd, want["nodes"]), (field, n, exchange, chunks)
                    for p, r, d in allrows:
                        assert np.array_equal(r.reshape(len(p), total_cols, -1), want["lde"][0][p][:, :total_cols].reshape(len(p), total_cols, -1))
                        assert np.array_equal(d, want["leaves"][p])
                h.close()
            assert roots[0] == roots[1]
        # the torch.distributed orchestration (parallel.py) over the same C-ABI device pieces
        be = DeviceBackend(field, hash_id, dev)
        for fused, pipeline in ((False, 0), (True, 0), (False, 2)):
            if fused and world > 8:
                continue
            dc = DistributedCommitter(be, n, blowup, total_cols, fused=fused, pipeline=pipeline)
            root = dc.commit(cols)
            nodes = dc.gather_nodes().numpy()
            if rank == 0:
                assert root == bytes(want["root"]) and np.array_equal(nodes, want["nodes"])
            dc.close()
    dist.barrier()
    if rank == 0:
        print("DISTRIBUTED_OK", flush=True)


def _horner(coeffs, x, p):
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * x + c) % p
    return acc


def cfg5(dev, log_rows, samples):
    import blake3

    import bench
    from starkpack_winterfell_b200 import sharded

    rank, world = dist.get_rank(), dist.get_world_size()
    n, C_, B = 1 << log_rows, 256, 8
    L = n * B
    s0, cnt = sharded.shard_columns(C_, world, rank)
    cols = bench.synth_trace_device(torch, cnt, n, bench.SEED5, dev, col_start=s0)
    polys = torch.empty_like(cols)
    t0 = time.perf_counter()
    h = sharded.commit_dev(0, 0, n, B, C_, cols, polys)
    torch.cuda.synchronize()
    first_ms = 1e3 * (time.perf_counter() - t0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h.close()
    dist.barrier()
    e0.record()
    h = sharded.commit_dev(0, 0, n, B, C_, cols, polys)
    e1.record()
    torch.cuda.synchronize()
    tms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    phases, sent = h.phases_ms()
    root = h.root()
    roots = [None] * world
    dist.all_gather_object(roots, root)
    assert all(r == root for r in roots), "ranks disagree on the root"
    g_lde = pow(7277203076849721926, 1 << (32 - (log_rows + 3)), P64)
    g_tr = pow(g_lde, B, P64)
    rng = np.random.default_rng(5)  # same stream on every rank
    checked = {"polys_back_to_trace": 0, "lde_entries_cross_rank": 0, "leaves_blake3_2chunks": 0, "merkle_paths": 0}
    # (a) + (b): sampled (column, LDE row) pairs; the column's owner evaluates P_c by Horner from ITS coefficients, the
    # row's owner reads the entry from ITS row block (which came through the exchange unless both owners coincide)
    pairs = [(int(rng.integers(0, C_)), int(rng.integers(0, L))) for _ in range(samples)]
    expected = {}
    for c, k in pairs:
        if s0 <= c < s0 + cnt:
            coeffs = [int(x) * RINV % P64 for x in polys[c - s0, :, 0].cpu().numpy().view(np.uint64)]
            r = k % n
            want_tr = int(cols[c - s0, r, 0].cpu().numpy().view(np.uint64)) * RINV % P64
            assert _horner(coeffs, pow(g_tr, r, P64), P64) == want_tr, "poly does not evaluate back to the trace"
            checked["polys_back_to_trace"] += 1
            expected[(c, k)] = _horner(coeffs, 7 * pow(g_lde, k, P64) % P64, P64)
    allexp = [None] * world
    dist.all_gather_object(allexp, expected)
    merged = {}
    for e in allexp:
        merged.update(e)
    for c, k in pairs:
        if h.row_start <= k < h.row_start + h.row_count:
            got = int(h.rows([k])[0, c]) * RINV % P64
            assert got == merged[(c, k)], f"LDE entry (col {c}, row {k}) differs from Horner"
            checked["lde_entries_cross_rank"] += 1
    # (c) + (d): leaves and Merkle paths for rows of this rank
    local_roots = [None] * world
    dist.all_gather_object(local_roots, bytes(h.digests(1, [1])[0]))
    H = lambda b: blake3.blake3(b).digest()  # noqa: E731
    for k in [h.row_start, h.row_start + h.row_count - 1] + [h.row_start + int(x) for x in rng.integers(0, h.row_count, 2)]:
        row = h.rows([k])[0]
        data = b"".join((int(x) * RINV % P64).to_bytes(8, "little") for x in row)
        assert len(data) == 2048
        kl = k - h.row_start
        leaf = bytes(h.digests(0, [kl])[0])
        assert H(data) == leaf, f"leaf {k} is not blake3 of the canonical row bytes"
        checked["leaves_blake3_2chunks"] += 1
        sib = bytes(h.digests(0, [kl ^ 1])[0])
        v = H(leaf + sib) if kl % 2 == 0 else H(sib + leaf)
        idx = (h.row_count + kl) >> 1
        while idx > 1:
            assert v == bytes(h.digests(1, [idx])[0])
            s = bytes(h.digests(1, [idx ^ 1])[0])
            v = H(v + s) if idx % 2 == 0 else H(s + v)
            idx >>= 1
        assert v == local_roots[rank]
        gi = world + rank  # the subtree root's index in the global heap
        while gi > 1:
            s = local_roots[(gi ^ 1) - world] if gi >= world else bytes(h.digests(2, [gi ^ 1])[0])
            v = H(v + s) if gi % 2 == 0 else H(s + v)
            gi >>= 1
            if gi >= 1 and gi < world:
                assert v == bytes(h.digests(2, [gi])[0])
        assert v == root
        checked["merkle_paths"] += 1
    tot = torch.tensor([checked[k] for k in sorted(checked)], device=dev)
    dist.all_reduce(tot)
    if rank == 0:
        rec = {"config": "cfg5", "rows": n, "cols": C_, "blowup": B, "n_gpus": world, "ms": float(tms.item()), "first_call_ms": first_ms,
               "exchange": h.exchange, "chunks": h.chunks, "phases_ms_rank0": phases, "bytes_sent_rank0": sent, "root": root.hex(),
               "checked": {k: int(v) for k, v in zip(sorted(checked), tot.tolist())}, "all_ranks_same_root": True}
        print(json.dumps(rec), flush=True)
    h.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg5", action="store_true")
    ap.add_argument("--log-rows", type=int, default=0)
    ap.add_argument("--samples", type=int, default=6)
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from starkpack_winterfell_b200 import sharded

    sharded.dist_init(dist.get_rank(), dist.get_world_size(), device=local)
    if args.cfg5:
        cfg5(dev, args.log_rows or (24 if dist.get_world_size() >= 4 else 22), args.samples)
    else:
        small_shapes(dev)
    sharded.dist_shutdown()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
