"""torchrun entry: ONE Merkle tree over a trace that is sharded by columns across G GPUs (parallel.DistributedCommitter:
per-rank interpolation + LDE, one NCCL all-to-all from column shards to row blocks, per-rank leaf hashing + subtree,
all-gather of the G subtree roots, top levels). Timed with CUDA events, max over ranks.
  cfg4: 8 fib traces of 2^20 x 2 f128, Blake3_256 (16 columns; at G = 8 one trace per GPU)
  cfg5: f64 Blake3_256, 2^LOG x 256 columns; LOG = 24 needs 8 GPUs (LDE shard + receive buffer = 69 GB per GPU)
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 \
         profiles/bench_distributed.py [--cfg5-log-rows 24]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import bench
from tests import fixtures as fx

ap = argparse.ArgumentParser()
ap.add_argument("--cfg5-log-rows", type=int, default=0, help="0: skip cfg5")
ap.add_argument("--cfg4-log-rows", type=int, default=20)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--pipeline", type=int, default=0, help="K > 1: LDE in K column chunks, chunk k's all-to-all under chunk k+1's LDE")
ap.add_argument("--auto", type=int, default=0, help="1: DistributedCommitter.auto picks the form of the exchange")
ap.add_argument("--fused", type=int, default=0, help="1: LDE stores rows into the owners' buffers over NVLink (no all-to-all)")
args = ap.parse_args()
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"
dist.init_process_group("nccl", device_id=dev)
from starkpack_winterfell_b200.parallel import DeviceBackend, DistributedCommitter  # noqa: E402

G, rank = dist.get_world_size(), dist.get_rank()
B = 8


def timed(dc, cols, name, n, total_cols, esize):
    for _ in range(2):
        root = dc.commit(cols)
    ts = []
    for _ in range(args.reps):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        root = dc.commit(cols)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    ts.sort()
    L = n * B
    alg = 2 * n * total_cols * esize + L * total_cols * esize + 2 * L * 32
    if rank == 0:
        print(json.dumps({"config": name, "n_gpus": G, "rows": n, "total_cols": total_cols, "gpu_ms_min": ts[0], "gpu_ms_median": ts[len(ts) // 2],
                          "lde_rows_per_s": L / (ts[0] * 1e-3), "algorithmic_bytes": alg, "hbm_gbs_per_gpu": alg / G / (ts[0] * 1e-3) / 1e9,
                          "exchange_bytes_sent_per_gpu": (L // G) * (G - 1) * ((dc.col_count + 7) // 8 * 8) * esize,
                          "exchange": "fused into the LDE pass (P2P stores)" if dc.fused else (f"NCCL all-to-all in {dc.pipeline} column chunks under the LDE" if dc.pipeline > 1 else "NCCL all-to-all"), "root": root.hex()}), flush=True)
    return root


# ---- cfg4
n = 1 << args.cfg4_log_rows
if 16 >= G:
    be = DeviceBackend(1, 0, dev)
    dc = DistributedCommitter.auto(be, n, B, 16) if args.auto else DistributedCommitter(be, n, B, 16, fused=bool(args.fused), pipeline=args.pipeline)
    traces = np.concatenate([fx.fib_trace_f128(n, start=(t + 1, t + 1)) for t in range(8)], axis=0)
    mine = np.ascontiguousarray(traces[dc.col_start: dc.col_start + dc.col_count]).view(np.int64).reshape(dc.col_count, n, 2)
    root = timed(dc, torch.from_numpy(mine).to(dev), "cfg4 starkpack 8 x fib f128 Blake3_256, column-sharded", n, 16, 16)
    if rank == 0 and n <= (1 << 20):
        import oracle

        oracle.build()
        want = oracle.commit_trace(1, 0, [fx.fib_trace_f128(n, start=(t + 1, t + 1)) for t in range(8)], B, want=("nodes",))
        assert bytes(want["root"]) == root, "cfg4 root differs from the oracle"
        print(json.dumps({"config": "cfg4", "matches_oracle": True}), flush=True)
    dc.close()
    del dc, be
    torch.cuda.empty_cache()

# ---- cfg5
if args.cfg5_log_rows:
    n = 1 << args.cfg5_log_rows
    C5 = 256
    be = DeviceBackend(0, 0, dev)
    dc = DistributedCommitter.auto(be, n, B, C5) if args.auto else DistributedCommitter(be, n, B, C5, fused=bool(args.fused), pipeline=args.pipeline)
    full_cols = bench.synth_trace_device(torch, dc.col_count, n, 0xC0FFEE05 + (dc.col_start << 32), dev)
    timed(dc, full_cols, f"cfg5 f64 2^{args.cfg5_log_rows} x 256 Blake3_256, column-sharded", n, C5, 8)
dist.barrier()
dist.destroy_process_group()
