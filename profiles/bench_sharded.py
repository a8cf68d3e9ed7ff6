"""torchrun entry: the headline trace (2^22 x 100 f64, blowup 8, Blake3_256; or --log-rows / --cols) committed through
wg_sharded_commit_dev with every form of the exchange, CUDA events, max over ranks; one JSON line per form with the
per-phase split (max over ranks) and the bytes each GPU sent. Used to choose WG_EXCHANGE_AUTO's policy.
usage: python -m torch.distributed.run --nproc-per-node N ... profiles/bench_sharded.py [--log-rows 22] [--cols 100] [--seed5]"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-rows", type=int, default=22)
    ap.add_argument("--cols", type=int, default=100)
    ap.add_argument("--forms", default="1:1,2:2,2:3,2:4,2:7,3:1,0:0")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--split-pref", type=int, default=0, help="wg_set_split_preferred_log (default 22; 24 = two-pass plan up to 2^24)")
    ap.add_argument("--split-log", type=int, default=0, help="wg_set_split_log: transforms above 2^k take the split plan (default 24)")
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    import starkpack_winterfell_b200 as w
    from starkpack_winterfell_b200 import sharded as sh

    sh.dist_init(rank, world, device=local)
    from starkpack_winterfell_b200._lib import lib
    if a.split_log:
        assert lib().wg_set_split_log(a.split_log) > 0
    if a.split_pref:
        assert lib().wg_set_split_preferred_log(a.split_pref) > 0
    n = 1 << a.log_rows
    seed = bench.SEED if a.cols == 100 else bench.SEED5
    s0, cnt = sh.shard_columns(a.cols, world, rank)
    cols = bench.synth_trace_device(torch, cnt, n, seed, dev, col_start=s0)
    roots = set()
    for form in a.forms.split(","):
        ex, k = (int(x) for x in form.split(":"))
        if ex == 3 and world > 8:
            continue
        best, ph_best, sent, used = None, None, 0, None
        for i in range(a.reps + 1):
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            h = sh.commit_dev(w.F64, w.BLAKE3_256, n, 8, a.cols, cols, exchange=ex, chunks=k)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ph, sent = h.phases_ms()
            names = list(ph)
            pt = torch.tensor([ph[x] for x in names], device=dev, dtype=torch.float64)
            dist.all_reduce(pt, op=dist.ReduceOp.MAX)
            if i > 0 and (best is None or float(t.item()) < best):
                best, ph_best = float(t.item()), {x: round(float(v), 3) for x, v in zip(names, pt.tolist())}
            used = (h.exchange, h.chunks)
            roots.add(h.root())
            h.close()
        if rank == 0:
            print(json.dumps({"n_gpus": world, "rows": n, "cols": a.cols, "asked": form, "split_log": a.split_log or 24, "split_pref": a.split_pref or 22, "exchange": used[0], "chunks": used[1], "ms": round(best, 3),
                              "phases_ms_max_over_ranks": ph_best, "bytes_sent_per_gpu": sent}), flush=True)
    assert len(roots) == 1, "exchange forms disagree on the root"
    if rank == 0:
        print(json.dumps({"root": roots.pop().hex(), "all_forms_same_root": True}), flush=True)
    sh.dist_shutdown()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
