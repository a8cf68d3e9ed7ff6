"""Times the non-headline BASELINE.json configs on ONE GPU, device-resident (CUDA events, 3 warm-ups, 10 timed commits,
min and median), checks each root against the CPU oracle run on the same input and times that oracle on the host cores.
  cfg1 fib      f128 / Blake3_256, 2^20 x 2
  cfg3 rescue   f128 / Sha3_256,   2^20 x 4  (uniform seed 0xC0FFEE03; SURVEY 8(d) allows it)
  cfg4 starkpack f128 / Blake3_256, 8 traces of 2^20 x 2 aggregated into one tree (here all 8 on one GPU)
usage: python profiles/bench_configs.py [--log-rows 20] > profiles/r1_configs.jsonl"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import oracle
import starkpack_winterfell_b200 as w
from starkpack_winterfell_b200.device import TraceCommitter
from tests import fixtures as fx

ap = argparse.ArgumentParser()
ap.add_argument("--log-rows", type=int, default=20)
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
n = 1 << args.log_rows
B = 8
oracle.build()


def algorithmic_bytes(n, cols, esize):
    L = n * B
    return 2 * n * cols * esize + L * cols * esize + 2 * L * 32


def run(name, field, hash_id, traces):
    esize = 8 if field == w.F64 else 16
    cpt = [t.shape[0] for t in traces]
    tc = TraceCommitter(field, hash_id, n, B, cpt, device="cuda:0")
    flat = np.concatenate(traces, axis=0)
    a = np.ascontiguousarray(flat).view(np.int64).reshape(flat.shape[0], n, esize // 8)
    tc.cols.copy_(torch.from_numpy(a))
    for _ in range(3):
        tc.commit()
    torch.cuda.synchronize()
    ts = []
    for _ in range(args.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tc.commit()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    root = tc.root()
    t0 = time.perf_counter()
    want = oracle.commit_trace(field, hash_id, traces, B, want=("nodes",))
    cpu_ms = 1e3 * (time.perf_counter() - t0)
    ok = bytes(want["root"]) == root and np.array_equal(tc.nodes.cpu().numpy(), want["nodes"])
    alg = algorithmic_bytes(n, sum(cpt), esize)
    print(json.dumps({"config": name, "rows": n, "cols_per_trace": cpt, "blowup": B, "gpu_ms_min": ts[0], "gpu_ms_median": ts[len(ts) // 2],
                      "lde_rows_per_s": n * B / (ts[0] * 1e-3), "algorithmic_bytes": alg, "hbm_gbs": alg / (ts[0] * 1e-3) / 1e9,
                      "cpu_oracle_ms": cpu_ms, "cpu_threads": int(oracle.lib().or_threads()), "root": root.hex(),
                      "matches_oracle": bool(ok)}), flush=True)
    assert ok, name
    del tc
    torch.cuda.empty_cache()


run("cfg1 fib f128 Blake3_256", w.F128, w.BLAKE3_256, [fx.fib_trace_f128(n)])
run("cfg3 rescue-shape f128 Sha3_256", w.F128, w.SHA3_256, [fx.synth_f128(0xC0FFEE03, 4, n)])
run("cfg4 starkpack 8 x fib f128 Blake3_256 (one GPU)", w.F128, w.BLAKE3_256, [fx.fib_trace_f128(n, start=(t + 1, t + 1)) for t in range(8)])
