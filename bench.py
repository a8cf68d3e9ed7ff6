#!/usr/bin/env python
"""bench.py — trace LDE + Merkle commit (BASELINE.json metric) on N B200s of one node.

One "step" = one full Prover::build_trace_commitment pass (interpolate columns -> coset LDE -> hash LDE rows ->
Merkle tree) over ONE synthetic trace = BASELINE.json configs[1]: 2^22 rows x 100 Goldilocks columns, blowup 8,
Blake3_256.

  N = 1   value  device-resident: trace already in HBM, results left in HBM, CUDA events around K steps
  N > 1   value  the SAME trace (same seed on every rank), its columns cut into N contiguous blocks, committed through
                 the C ABI's wg_sharded_commit_dev: per-rank interpolation + LDE, the column-shard -> row-block exchange
                 over NVLink (NCCL all-to-all pipelined under the LDE, or P2P stores fused into the LDE's last pass), row
                 hashing, the subtree, the root all-gather and the top levels. "strong" scaling: value = ms per commit,
                 CUDA events, max over ranks. The root is asserted equal to the 1-GPU root on every N.
  e2e          the same work through the C-ABI host entry point with pinned host buffers: H2D of the trace and D2H of
               polys + LDE + leaves + nodes inside the timed region (N > 1: each rank moves its column block in and its
               row block out)
  roofline     dominant kernel (LDE contiguous pass): SURVEY 8(d) algorithmic bytes it moves / its CUDA-event time
  cpu_baseline the CPU oracle (C + OpenMP restatement of the reference's `concurrent` path, kind "port") on a bounded sample
  replicas     (N > 1) the old figure: one independent trace per GPU, no exchange
  configs      cfg4 (8 fib traces, one tree) and, where it fits, cfg5 (2^24 x 256) through the same sharded entry point
  --impl reference : that oracle as the reference arm, on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG_ROWS, COLS, BLOWUP = 22, 100, 8
SEED = 0xC0FFEE02
SEED5 = 0xC0FFEE05
METRIC = "lde_merkle_commit_ms_2^22x100_f64_blowup8"


def golden_roots():
    """full-size roots from the CPU oracle (tests/golden/make_full_roots.py); {} if the file is absent"""
    try:
        with open(os.path.join(ROOT, "tests", "golden", "full_roots.json")) as f:
            return {k: v["root"] for k, v in json.load(f).items()}
    except Exception:
        return {}


def algorithmic_bytes(n, cols, blowup, esize=8):
    """read trace + write polys + write LDE (unpadded) + write leaves + write nodes."""
    L = n * blowup
    return n * cols * esize * 2 + L * cols * esize + 2 * L * 32


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def synth_trace_device(torch, cols, n, seed, device, col_start=0):
    """SURVEY 8(d) generator x(c, r) = splitmix64(seed + c*2^32 + r) reduced below p, computed on the GPU with
    int64 wraparound arithmetic (identical words to tests/fixtures.synth_f64); columns col_start .. col_start+cols."""
    def lsr(x, k):
        return (x >> k) & ((1 << (64 - k)) - 1)

    def i64(v):
        v &= (1 << 64) - 1
        return v - (1 << 64) if v >> 63 else v

    out = torch.empty((cols, n, 1), dtype=torch.int64, device=device)
    r = torch.arange(n, dtype=torch.int64, device=device)
    for c in range(cols):
        z = r + i64(seed + ((c + col_start) << 32)) + i64(0x9E3779B97F4A7C15)
        z = (z ^ lsr(z, 30)) * i64(0xBF58476D1CE4E5B9)
        z = (z ^ lsr(z, 27)) * i64(0x94D049BB133111EB)
        z = z ^ lsr(z, 31)
        # unsigned z >= p  <=>  signed z in [p - 2^64, -1] = [-(2^32) + 1, -1]; subtracting p == adding 2^32 - 1
        ge = (z < 0) & (z >= -(1 << 32) + 1)
        out[c, :, 0] = torch.where(ge, z + ((1 << 32) - 1), z)
    return out


# ---------------------------------------------------------------------------------------------------------- CPU arm
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def pick_cpu_log_rows(requested):
    """Full 2^22 rows when the host can do it in bounded time (>= 16 threads, >= 96 GB free: ~60 GB working set, 50 - 80 s),
    otherwise a 2^20-row sample scaled linearly in rows."""
    if requested:
        return requested
    try:
        import psutil

        free = psutil.virtual_memory().available
    except Exception:
        free = 0
    return LOG_ROWS if (host_threads() >= 16 and free > 96e9) else 20


def cpu_commit_sample(log_rows_sample, runs=1):
    """Times the CPU oracle (oracle/oracle.c: C + OpenMP restatement of the reference's `concurrent` rayon path — nested
    parallel coset / row FFTs, concurrent permute, batched transpose, parallel leaf hashing and Merkle levels; kind "port":
    no Rust toolchain here, so oracle/_ref does not exist) with every host thread. torchrun exports OMP_NUM_THREADS=1, so
    the thread count is set explicitly."""
    import oracle
    from tests import fixtures as fx

    oracle.build()
    threads = host_threads()
    oracle.lib().or_set_threads(threads)
    n = 1 << log_rows_sample
    tr = fx.synth_f64(SEED, COLS, n)
    times, root = [], None
    for _ in range(max(1, runs)):
        t0 = time.perf_counter()
        r = oracle.commit_trace(0, 0, [tr], BLOWUP, want=("nodes",))
        times.append(time.perf_counter() - t0)
        root = bytes(r["root"])
    ms_sample = 1e3 * min(times)
    scale = (1 << LOG_ROWS) / n
    what = "the full workload" if scale == 1 else f"scaled x{scale:g} (linear in rows) to 2^{LOG_ROWS}"
    return {"value": ms_sample * scale, "unit": "ms", "cores": int(oracle.lib().or_threads()), "kind": "port",
            "sample": f"2^{log_rows_sample} rows x {COLS} cols f64 blowup {BLOWUP} Blake3_256, {ms_sample:.1f} ms measured "
                      f"(best of {len(times)}), {what}",
            "sample_ms": ms_sample, "sample_root": root.hex(), "runs": len(times), "host_threads": threads}


def workload_config(world):
    part = "one GPU" if world == 1 else (f"columns sharded over {world} GPUs (wg_sharded_commit_dev: LDE per column block, row-block "
                                         "exchange over NVLink, row hashing + subtree per GPU, top levels)")
    return {"workload": f"2^{LOG_ROWS} rows x {COLS} cols f64 (Goldilocks), blowup {BLOWUP}, Blake3_256 commit of ONE trace, {part}",
            "l2": "inputs (3.4 GB trace, 27 GB LDE) far exceed the 126 MB L2; no flush needed",
            "value_is": "ms per commit of the one trace (max over ranks)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    log_rows = pick_cpu_log_rows(args.cpu_log_rows)
    runs = 1 if log_rows >= 22 else min(3, max(1, args.steps))  # bounded: the sample is never repeated `steps` times
    cb = cpu_commit_sample(log_rows, runs=runs)
    gold = golden_roots().get("cfg2")
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "ms", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["sample_ms"], "higher_is_better": False,
            "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": workload_config(world),
            "reference_arm": "CPU oracle port of the reference's concurrent path on the host cores (Rust toolchain absent: no "
                             f"oracle/_ref); {cb['cores']} OpenMP threads; the sample is run {cb['runs']}x whatever --steps says",
            "cpu_baseline": cb,
            "root_matches_golden": (cb["sample_root"] == gold) if (gold and log_rows == LOG_ROWS) else None,
            "e2e": {"value": cb["value"], "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------- e2e legs
def _max_over_ranks(torch, dist, dev, ms):
    if dist is None:
        return ms
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_e2e(torch, w, dev, steps, warmup):
    """N = 1: wg_commit_trace through host buffers: pinned trace in, pinned polys + LDE + leaves + nodes out."""
    import ctypes as C

    from starkpack_winterfell_b200._lib import check, lib

    n, L = 1 << LOG_ROWS, (1 << LOG_ROWS) * BLOWUP
    rw = (COLS + 7) // 8 * 8
    pin = dict(dtype=torch.int64, pin_memory=True)
    cols_h = torch.empty((COLS, n), **pin)
    cols_h.copy_(synth_trace_device(torch, COLS, n, SEED, dev)[:, :, 0])
    polys_h = torch.empty((COLS, n), **pin)
    lde_h = torch.empty((L, rw), **pin)
    leaves_h = torch.empty((L, 32), dtype=torch.uint8, pin_memory=True)
    nodes_h = torch.empty((L, 32), dtype=torch.uint8, pin_memory=True)
    root_h = torch.zeros(32, dtype=torch.uint8, pin_memory=True)
    torch.cuda.synchronize()
    colp = (C.c_void_p * COLS)(*[cols_h[c].data_ptr() for c in range(COLS)])
    polp = (C.c_void_p * COLS)(*[polys_h[c].data_ptr() for c in range(COLS)])
    ldep = (C.c_void_p * 1)(lde_h.data_ptr())
    cpt = (C.c_uint32 * 1)(COLS)

    def step():
        check(lib().wg_commit_trace(w.F64, w.BLAKE3_256, n, BLOWUP, 1, cpt, colp, polp, ldep, C.c_void_p(leaves_h.data_ptr()),
                                    C.c_void_p(nodes_h.data_ptr()), C.c_void_p(root_h.data_ptr())))

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()  # synchronous: returns after the D2H copies have landed
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0) / steps
    h2d = n * COLS * 8
    d2h = n * COLS * 8 + L * rw * 8 + 2 * L * 32 + 32
    return {"value": ms, "unit": "ms", "ms_per_step": ms, "steps": steps, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h, "root": bytes(root_h.numpy()).hex(), "outputs": "polys + LDE + leaves + nodes + root",
            "timer": "host perf_counter around synchronous C-ABI calls (wg_commit_trace)"}


def run_e2e_resident(torch, w, dev, steps, warmup):
    """N = 1: the resident C-ABI entry (wg_commitment_from_trace): pinned trace in, pinned polys + root out, then
    TraceCommitment::query for 32 positions (rows + batch Merkle proof); LDE and tree stay in HBM."""
    import numpy as np

    from starkpack_winterfell_b200.matrix import ColMatrix, StarkDomain
    from starkpack_winterfell_b200.prover import ResidentCommitment

    n = 1 << LOG_ROWS
    cols_h = torch.empty((COLS, n), dtype=torch.int64, pin_memory=True)
    cols_h.copy_(synth_trace_device(torch, COLS, n, SEED, dev)[:, :, 0])
    torch.cuda.synchronize()
    cols_np = cols_h.numpy().view(np.uint64)
    trace = ColMatrix(w.F64, [cols_np[c] for c in range(COLS)])
    dom = StarkDomain(w.F64, n, BLOWUP)
    polys_pinned = torch.empty((COLS, n), dtype=torch.int64, pin_memory=True)
    polys_np = polys_pinned.numpy().view(np.uint64)
    polys_out = [[polys_np[c] for c in range(COLS)]]
    rng = np.random.default_rng(1)
    pos = [int(x) for x in rng.choice(n * BLOWUP, size=32, replace=False)]
    root = None
    d2h = 0

    def step():
        nonlocal root, d2h
        rc = ResidentCommitment.from_traces(w.BLAKE3_256, [trace], dom, polys_out=polys_out)
        q = rc.query(pos)
        root = rc.root()
        d2h = n * COLS * 8 + 32 + q.traces_states[0].nbytes + 32 * (len(q.trace_proof.leaves) + sum(len(v) for v in q.trace_proof.nodes))
        rc.close()

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0) / steps
    return {"value": ms, "unit": "ms", "ms_per_step": ms, "steps": steps, "h2d_bytes_per_step": n * COLS * 8,
            "d2h_bytes_per_step": int(d2h), "root": root.hex(),
            "outputs": "polys + root + 32 queried rows with their batch Merkle proof; LDE and tree stay in HBM "
                       "(wg_commitment_from_trace / _rows / _digests / _free)"}


def run_e2e_sharded(torch, w, sh, dev, rank, world, steps, warmup, dist):
    """N > 1: wg_sharded_commit through host buffers. Every rank uploads ITS column block from pinned memory, the library
    commits, and every rank downloads its polynomials, its row block of the LDE (all columns of rows [rank*L/N, ...)),
    its leaves and subtree and the top levels: together exactly the outputs of the 1-GPU call, each moved once."""
    import ctypes as C

    import numpy as np

    from starkpack_winterfell_b200._lib import check, lib

    n = 1 << LOG_ROWS
    s0, cnt = sh.shard_columns(COLS, world, rank)
    cols_h = torch.empty((cnt, n), dtype=torch.int64, pin_memory=True)
    cols_h.copy_(synth_trace_device(torch, cnt, n, SEED, dev, col_start=s0)[:, :, 0])
    polys_h = torch.empty((cnt, n), dtype=torch.int64, pin_memory=True)
    torch.cuda.synchronize()
    cols_np, polys_np = cols_h.numpy().view(np.uint64), polys_h.numpy().view(np.uint64)
    cols_list, polys_list = [cols_np[c] for c in range(cnt)], [polys_np[c] for c in range(cnt)]
    bufs = {}
    root, d2h = None, 0

    def step():
        nonlocal root, d2h
        h = sh.commit_host(w.F64, w.BLAKE3_256, n, BLOWUP, COLS, cols_list, polys_out=polys_list)
        if not bufs:
            bufs["blocks"] = [torch.empty((h.row_count, rw), dtype=torch.int64, pin_memory=True) if c else None
                              for c, rw in zip(h.block_cols, h.block_row_widths)]
            bufs["leaves"] = torch.empty((h.row_count, 32), dtype=torch.uint8, pin_memory=True)
            bufs["nodes"] = torch.empty((h.row_count, 32), dtype=torch.uint8, pin_memory=True)
            bufs["top"] = torch.empty((max(world, 2), 32), dtype=torch.uint8, pin_memory=True)
        ptrs = (C.c_void_p * len(bufs["blocks"]))(*[b.data_ptr() if b is not None else None for b in bufs["blocks"]])
        check(lib().wg_sharded_download(h._h, ptrs, C.c_void_p(bufs["leaves"].data_ptr()), C.c_void_p(bufs["nodes"].data_ptr()),
                                        C.c_void_p(bufs["top"].data_ptr())))
        root = h.root()
        d2h = cnt * n * 8 + sum(b.numel() * 8 for b in bufs["blocks"] if b is not None) + 2 * h.row_count * 32 + max(world, 2) * 32
        h.close()

    for _ in range(warmup):
        step()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    ms = _max_over_ranks(torch, dist, dev, 1e3 * (time.perf_counter() - t0) / steps)
    tot = torch.tensor([cnt * n * 8, d2h], device=dev, dtype=torch.float64)
    dist.all_reduce(tot)
    return {"value": ms, "unit": "ms", "ms_per_step": ms, "steps": steps, "h2d_bytes_per_step": int(tot[0].item()),
            "d2h_bytes_per_step": int(tot[1].item()), "root": root.hex(),
            "outputs": "per rank: its polys + its row block of the LDE (all columns) + its leaves + subtree + top levels; bytes summed over ranks",
            "timer": "host perf_counter around synchronous C-ABI calls (wg_sharded_commit + wg_sharded_download), barrier on both sides, max over ranks"}


# ---------------------------------------------------------------------------------------------------------- other configs
def fib_trace_words(torch, n, start, dev):
    """prover/src/tests/mod.rs:17-29 recurrence mod q as [2, n, 2] int64 (lo, hi) words; Python ints on the host."""
    import numpy as np

    q = (1 << 128) - 45 * (1 << 40) + 1
    a, b = start
    lo = np.empty((2, n), dtype=np.uint64)
    hi = np.empty((2, n), dtype=np.uint64)
    m = (1 << 64) - 1
    la, ha, lb, hb = [], [], [], []
    for _ in range(n):
        la.append(a & m)
        ha.append(a >> 64)
        lb.append(b & m)
        hb.append(b >> 64)
        a = (a + b) % q
        b = (a + b) % q
    lo[0], hi[0], lo[1], hi[1] = la, ha, lb, hb
    out = np.stack([lo, hi], axis=-1)
    return torch.from_numpy(out.view(np.int64)).to(dev)


def run_config_sharded(torch, w, sh, dist, dev, rank, world, name, field, hash_id, n, total_cols, make_local, reps=3):
    """One of the other BASELINE.json configs through wg_sharded_commit_dev; CUDA events, max over ranks, best of reps."""
    s0, cnt = sh.shard_columns(total_cols, world, rank)
    local = make_local(s0, cnt)
    torch.cuda.synchronize()
    best, root, ph, sent, ex, k = None, None, None, 0, None, 0
    for i in range(reps + 1):
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h = sh.commit_dev(field, hash_id, n, BLOWUP, total_cols, local)
        e1.record()
        torch.cuda.synchronize()
        ms = _max_over_ranks(torch, dist, dev, e0.elapsed_time(e1))
        if i > 0 and (best is None or ms < best):
            best = ms
            ph, sent = h.phases_ms()
        root, ex, k = h.root(), h.exchange, h.chunks
        h.close()
    del local
    torch.cuda.empty_cache()
    esize = 8 if field == w.F64 else 16
    alg = algorithmic_bytes(n, total_cols, BLOWUP, esize)
    gbs = alg / world / (best * 1e-3) / 1e9
    return {"config": name, "rows": n, "total_cols": total_cols, "n_gpus": world, "ms": best, "exchange": ex, "chunks": k,
            "phases_ms_rank0": {a: round(b, 3) for a, b in ph.items()}, "bytes_sent_rank0": sent, "root": root.hex(),
            "lde_rows_per_s": n * BLOWUP / (best * 1e-3), "trace_rows_per_s": n / (best * 1e-3),
            "algorithmic_bytes": alg, "hbm_gbs_per_gpu": gbs, "hbm_frac_per_gpu": gbs / peaks()[0]}


# ---------------------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    # stdout carries exactly ONE JSON line: anything libraries print meanwhile (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        import torch.distributed as dist_mod

        dist_mod.init_process_group("nccl", device_id=dev)
        dist = dist_mod
    import starkpack_winterfell_b200 as w
    from starkpack_winterfell_b200 import _lib
    from starkpack_winterfell_b200 import sharded as sh
    from starkpack_winterfell_b200.device import TraceCommitter
    from starkpack_winterfell_b200.hostmem import bind_to_gpu_node

    sh.dist_init(rank, world, device=local)  # wg_init + (N > 1) the library's own NCCL communicator
    gold = golden_roots()
    n = 1 << LOG_ROWS
    L = n * BLOWUP
    W, K = max(args.warmup, 3), max(args.steps, 1)
    sampler = ClockSampler(local)
    extra = {}

    if world == 1:
        tc = TraceCommitter(w.F64, w.BLAKE3_256, n, BLOWUP, [COLS], device=dev)
        tc.cols.copy_(synth_trace_device(torch, COLS, n, SEED, dev))
        for _ in range(W):
            tc.commit()
        torch.cuda.synchronize()
        root_warm = tc.root()
        _lib.profile_enable(True)
        _lib.profile_read()
        sampler.start()
        launches0 = _lib.launch_count()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
        torch.cuda.synchronize()
        evs[0].record()
        for i in range(K):
            tc.commit()
            evs[i + 1].record()
        torch.cuda.synchronize()
        ms = evs[0].elapsed_time(evs[K]) / K
        per_step = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
        launches = _lib.launch_count() - launches0
        clocks = sampler.stop()
        phases = {k: (v[0] / K, v[1] // K) for k, v in _lib.profile_read().items()}
        _lib.profile_enable(False)
        assert tc.root() == root_warm, "root changed between identical steps"
        root_hex = tc.root().hex()
        del tc
        torch.cuda.empty_cache()
        _lib.check(_lib.lib().wg_shutdown())
        # pinned host buffers on the GPU's NUMA node (hostmem.py): the host-buffer legs are PCIe-bound
        with bind_to_gpu_node(local) as numa:
            e2e = run_e2e(torch, w, dev, steps=min(K, 2), warmup=1)
            torch.cuda.empty_cache()
            extra["e2e_resident"] = run_e2e_resident(torch, w, dev, steps=min(K, 2), warmup=1)
        e2e["host_numa"] = numa.info
        assert extra["e2e_resident"]["root"] == root_hex == e2e["root"], "resident / host-pointer / device-resident roots differ"
    else:
        s0, cnt = sh.shard_columns(COLS, world, rank)
        local_cols = synth_trace_device(torch, cnt, n, SEED, dev, col_start=s0)
        root_hex = None
        for _ in range(W):
            h = sh.commit_dev(w.F64, w.BLAKE3_256, n, BLOWUP, COLS, local_cols)
            root_hex = h.root().hex()
            h.close()
        torch.cuda.synchronize()
        _lib.profile_enable(True)
        _lib.profile_read()
        if rank == 0:
            sampler.start()
        launches0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        evs[0].record()
        shard_ph, sent, ex, chunks = {}, 0, None, 0
        for i in range(K):
            h = sh.commit_dev(w.F64, w.BLAKE3_256, n, BLOWUP, COLS, local_cols)
            evs[i + 1].record()
            ph, sent = h.phases_ms()
            for a, b in ph.items():
                shard_ph[a] = shard_ph.get(a, 0.0) + b / K
            assert h.root().hex() == root_hex, "root changed between identical steps"
            ex, chunks = h.exchange, h.chunks
            h.close()
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = _max_over_ranks(torch, dist, dev, e0.elapsed_time(e1) / K)
        ps = torch.tensor([evs[i].elapsed_time(evs[i + 1]) for i in range(K)], device=dev, dtype=torch.float64)
        dist.all_reduce(ps, op=dist.ReduceOp.MAX)
        per_step = ps.tolist()
        launches = _lib.launch_count() - launches0
        clocks = sampler.stop() if rank == 0 else None
        phases = {k: (v[0] / K, v[1] // K) for k, v in _lib.profile_read().items()}
        _lib.profile_enable(False)
        # per-phase maxima over ranks name the limiting piece
        names = list(shard_ph)
        t = torch.tensor([shard_ph[a] for a in names], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        exposed = shard_ph.get("exchange_exposed", 0.0)
        extra["sharded"] = {"exchange": ex, "chunks": chunks, "cols_rank0": cnt, "bytes_sent_per_gpu": sent,
                            "phases_ms_max_over_ranks": {a: round(float(v), 3) for a, v in zip(names, t.tolist())},
                            "nvlink_gbs_per_gpu_exposed_part_only": (sent / (exposed * 1e-3) / 1e9) if exposed > 0.05 else None}
        del local_cols
        torch.cuda.empty_cache()
        _lib.check(_lib.lib().wg_shutdown())
        with bind_to_gpu_node(local) as numa:  # every rank's pinned buffers on ITS GPU's NUMA node
            e2e = run_e2e_sharded(torch, w, sh, dev, rank, world, steps=min(K, 2), warmup=1, dist=dist)
        e2e["host_numa_rank0"] = numa.info
        assert e2e["root"] == root_hex, "host-pointer sharded root differs"
        torch.cuda.empty_cache()
        _lib.check(_lib.lib().wg_shutdown())
        # the old figure: one independent trace per GPU, no exchange (replicas)
        tc = TraceCommitter(w.F64, w.BLAKE3_256, n, BLOWUP, [COLS], device=dev)
        tc.cols.copy_(synth_trace_device(torch, COLS, n, SEED, dev))
        tc.commit()
        dist.barrier()
        torch.cuda.synchronize()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        for _ in range(2):
            tc.commit()
        r1.record()
        torch.cuda.synchronize()
        rep_ms = _max_over_ranks(torch, dist, dev, r0.elapsed_time(r1) / 2)
        extra["replicas"] = {"ms_per_step": rep_ms, "ms_per_trace": rep_ms / world, "traces_per_step": world, "root_rank0": tc.root().hex(),
                             "note": "N independent 2^22 x 100 commits, no data-path collective (round 1's `value`)"}
        assert tc.root().hex() == root_hex, "sharded root differs from this GPU's own single-GPU commit of the same trace"
        del tc
        torch.cuda.empty_cache()
        _lib.check(_lib.lib().wg_shutdown())

    if gold.get("cfg2"):
        assert root_hex == gold["cfg2"], f"root {root_hex} differs from the CPU oracle's full-size root {gold['cfg2']}"

    # ---- the other multi-GPU configs of BASELINE.json through the same entry point
    configs = []
    if not args.no_configs:
        n20 = 1 << 20

        def cfg4_local(s0, cnt):
            parts = []
            for t in range(s0 // 2, (s0 + cnt + 1) // 2):
                tr = fib_trace_words(torch, n20, (t + 1, t + 1), dev)
                lo, hi = max(s0, 2 * t), min(s0 + cnt, 2 * t + 2)
                parts.append(tr[lo - 2 * t: hi - 2 * t])
            return torch.cat(parts).contiguous()

        c4 = run_config_sharded(torch, w, sh, dist, dev, rank, world, "cfg4 starkpack: 8 fib traces f128 2^20 x 2, one tree", w.F128,
                                w.BLAKE3_256, n20, 16, cfg4_local)
        c4["matches_oracle_root"] = (c4["root"] == gold["cfg4"]) if gold.get("cfg4") else None
        configs.append(c4)
        if world == 1:
            # cfg1 / cfg3: one narrow f128 trace each (2 and 4 columns): they do not shard, so they are timed on one GPU only
            c1 = run_config_sharded(torch, w, sh, dist, dev, rank, world, "cfg1 fib: f128 2^20 x 2, Blake3_256", w.F128, w.BLAKE3_256, n20, 2,
                                    lambda s0, cnt: fib_trace_words(torch, n20, (1, 1), dev)[s0:s0 + cnt].contiguous())
            c1["matches_oracle_root"] = (c1["root"] == gold["cfg1"]) if gold.get("cfg1") else None
            configs.append(c1)
            from tests import fixtures as fx
            import numpy as np

            t3 = torch.from_numpy(np.ascontiguousarray(fx.synth_f128(0xC0FFEE03, 4, n20)).view(np.int64)).to(dev)
            c3 = run_config_sharded(torch, w, sh, dist, dev, rank, world, "cfg3 rescue shape: f128 2^20 x 4, Sha3_256", w.F128, w.SHA3_256, n20, 4,
                                    lambda s0, cnt: t3[s0:s0 + cnt].contiguous())
            c3["matches_oracle_root"] = (c3["root"] == gold["cfg3"]) if gold.get("cfg3") else None
            configs.append(c3)
            del t3
        else:
            configs.append({"config": "cfg1 fib / cfg3 rescue shape (one f128 trace of 2 / 4 columns)", "n_gpus": world,
                            "skipped": "a single narrow trace does not shard (fewer column tiles than GPUs): replicas only; timed at N = 1"})
        if world >= 4:
            n24 = 1 << 24
            c5 = run_config_sharded(torch, w, sh, dist, dev, rank, world, "cfg5 big: f64 2^24 x 256", w.F64, w.BLAKE3_256, n24, 256,
                                    lambda s0, cnt: synth_trace_device(torch, cnt, n24, SEED5, dev, col_start=s0), reps=2)
            c5["checked"] = "root equal across N; sampled LDE / leaf / path properties: tests/run_distributed_gpu.py --cfg5 (record under profiles/)"
            configs.append(c5)
        else:
            configs.append({"config": "cfg5 big: f64 2^24 x 256", "n_gpus": world, "skipped": "needs >= 4 GPUs: at N = 2 the 137 GB row block "
                            "per GPU does not fit next to the LDE send buffers (DESIGN.md)"})
        _lib.check(_lib.lib().wg_shutdown())

    sh.dist_shutdown()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    # Dominant kernel: the LDE contiguous pass (ntt_pass_kernel<Gl64,11,1,false>) and the LDE strided pass take within a
    # few percent of each other; the contiguous pass is the one that moves SURVEY 8(d)'s algorithmic bytes (it writes the
    # L x C LDE words, 26.8 of the 35.7 GB), so it is the one quoted. `achieved` counts ONLY those algorithmic bytes;
    # `kernel_io_*` also counts the kernel's own compulsory read of the B*N*C intermediate. N > 1: rank 0's column block.
    dom = "lde_contig"
    dom_by_time = max(phases, key=lambda k: phases[k][0])
    my_cols = COLS if world == 1 else sh.shard_columns(COLS, world, 0)[1]
    dom_launches = max(phases[dom][1], 1)
    dom_ms = phases[dom][0] / dom_launches
    kbytes_launch = L * my_cols * 8 / dom_launches
    kio_launch = 2 * kbytes_launch
    achieved = kbytes_launch / (dom_ms * 1e-3) / 1e9
    # cited constants from one ncu pass at this size (profiles/r2_traffic.json names the commit they were taken at): DRAM
    # bytes and executed warp-instructions per launch. int_issue_frac = those instructions / the kernel time measured in
    # THIS run / the measured integer peak (1.0 warp-instruction per clock per SMSP, profiles/r2_int_peak_summary.md).
    traffic, traffic_src, int_issue = None, None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            tj = json.load(f)
        if world == 1:
            traffic = tj["phases"][dom]["dram_bytes"]
            traffic_src = ("cited constant: ncu dram__bytes_read.sum + dram__bytes_write.sum of one launch at this size, captured at "
                           f"commit {tj.get('commit', '?')} (profiles/r2_traffic.json); not re-measured in this run")
            ipeak = float(tj.get("int_peak_warp_inst_per_s", 592 * 1.965e9))
            int_issue = {"peak_warp_inst_per_s": ipeak, "peak_source": "profiles/r2_int_peak_summary.md (measured: ALU pipe 0.5 + FMA pipe 0.5 "
                         "warp-instructions per clock per SMSP at 1965 MHz)", "instructions_source": traffic_src.replace(
                             "dram__bytes_read.sum + dram__bytes_write.sum", "smsp__inst_executed.sum"), "kernels": {}}
            for ph, (tms, nl) in phases.items():
                wi = tj["phases"].get(ph, {}).get("warp_instructions")
                if wi and tms > 0:
                    int_issue["kernels"][ph] = {"warp_instructions": wi, "frac": wi * max(nl, 1) / (tms * 1e-3) / ipeak}
            if dom in int_issue["kernels"]:
                int_issue["frac"] = int_issue["kernels"][dom]["frac"]
    except Exception:
        pass
    alg = algorithmic_bytes(n, COLS, BLOWUP)
    cb = cpu_commit_sample(pick_cpu_log_rows(args.cpu_log_rows)) if world == 1 else None
    if cb and "the full workload" in cb["sample"]:
        cb["root_matches_gpu"] = cb["sample_root"] == root_hex
    line = {
        "metric": METRIC, "value": ms, "unit": "ms", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms,
        "higher_is_better": False, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(world),
        "ms_min": min(per_step), "ms_median": sorted(per_step)[len(per_step) // 2],
        "lde_rows_per_s": L / (ms * 1e-3), "trace_rows_per_s": n / (ms * 1e-3),
        "commit_roofline": {"algorithmic_bytes": alg, "achieved_gbs_per_gpu": alg / world / (ms * 1e-3) / 1e9,
                            "frac_per_gpu": alg / world / (ms * 1e-3) / 1e9 / peak},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "root": root_hex,
        "root_matches_cpu_oracle_full_size": (root_hex == gold["cfg2"]) if gold.get("cfg2") else None,
        "phases_ms": {k: round(v[0], 3) for k, v in phases.items()},
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "int_issue_frac": (int_issue or {}).get("frac"), "int_issue": int_issue, "algorithmic_bytes_per_launch": kbytes_launch,
                     "kernel_ms": dom_ms, "launches_per_step": dom_launches,
                     "kernel_share_of_step": phases[dom][0] / ms, "longest_phase": dom_by_time,
                     "kernel_io_bytes": kio_launch, "kernel_io_gbs": kio_launch / (dom_ms * 1e-3) / 1e9,
                     "kernel_io_frac": kio_launch / (dom_ms * 1e-3) / 1e9 / peak,
                     "note": "integer-issue bound (ALU and FMA pipes both ~60% busy, DRAM ~25%), see DESIGN.md and "
                             "profiles/r2_int_peak.jsonl; HBM fraction reported as the contract asks"},
        "cpu_baseline": cb, "configs": configs,
    }
    line.update(extra)
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-log-rows", type=int, default=0, help="rows (log2) of the CPU baseline sample; 0 = 22 when the host has "
                    ">= 16 threads and >= 96 GB free, else 20")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg4 / cfg5 sub-benchmarks")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
