"""Full-size roots of the BASELINE.json configs from the CPU oracle (oracle/oracle.c), committed as
tests/golden/full_roots.json so that the GPU tests and bench.py can assert the full-size commits against a value the
CUDA path did not produce. Run on any host with ~48 GB of RAM:  python tests/golden/make_full_roots.py [cfg ...]
  cfg1  fib f128 2^20 x 2, Blake3_256          cfg3  f128 2^20 x 4 (seed 0xC0FFEE03), Sha3_256
  cfg2  f64 2^22 x 100 (seed 0xC0FFEE02)       cfg4  8 fib traces (start (t+1, t+1)) f128 2^20 x 2, one tree
cfg5 (2^24 x 256: 275 GB of LDE) does not fit a host; it is checked by sampled properties instead
(tests/run_distributed_gpu.py --cfg5)."""
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np

import oracle
from tests import fixtures as fx

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "full_roots.json")


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run(name, field, hash_id, traces, blowup=8):
    t0 = time.perf_counter()
    r = oracle.commit_trace(field, hash_id, traces, blowup, want=("leaves", "nodes"))
    dt = time.perf_counter() - t0
    return {"root": bytes(r["root"]).hex(), "leaves_sha256": digest(r["leaves"]), "nodes_sha256": digest(r["nodes"]),
            "rows": int(traces[0].shape[1]), "cols_per_trace": [int(t.shape[0]) for t in traces], "blowup": blowup,
            "field": field, "hash": hash_id, "oracle_seconds": round(dt, 1), "oracle_threads": int(oracle.lib().or_threads())}


def main():
    oracle.build()
    which = sys.argv[1:] or ["cfg1", "cfg3", "cfg4", "cfg2"]
    try:
        with open(OUT) as f:
            out = json.load(f)
    except Exception:
        out = {}
    n20 = 1 << 20
    for cfg in which:
        if cfg == "cfg1":
            out[cfg] = run(cfg, 1, 0, [fx.fib_trace_f128(n20)])
        elif cfg == "cfg3":
            out[cfg] = run(cfg, 1, 1, [fx.synth_f128(0xC0FFEE03, 4, n20)])
        elif cfg == "cfg4":
            out[cfg] = run(cfg, 1, 0, [fx.fib_trace_f128(n20, start=(t + 1, t + 1)) for t in range(8)])
        elif cfg == "cfg2":
            # the synthetic words (< p) ARE the trace's internal (Montgomery) words, exactly as bench.py uploads them
            out[cfg] = run(cfg, 0, 0, [fx.synth_f64(0xC0FFEE02, 100, 1 << 22)])
        print(cfg, out[cfg], flush=True)
        with open(OUT, "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
