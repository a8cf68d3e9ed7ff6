"""FRI commit phase at the examples' options (examples/src/lib.rs:97-104: blowup 8, folding 4,
remainder degree 31, FieldExtension::None) on BASELINE config 1's LDE domain (2^23 f128, Blake3_256) and on a 2^25 f64
domain (config 2's LDE rows). Prints one JSON line per case: GPU ms for FriProver.build_layers with the evaluations already
in HBM (CUDA events; includes the per-layer 32-byte root read the Fiat-Shamir channel needs), the CPU oracle beside it
(single thread: the reference's apply_drp is rayon-parallel, ours is a scalar restatement), roots compared.

    python profiles/bench_fri.py [--log-n 23] [--reps 5]
"""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import fixtures as fx  # noqa: E402
from tests.test_gpu_fri import HashChannel  # noqa: E402

F64, F128, B3 = 0, 1, 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cpu", type=int, default=1)
    a = ap.parse_args()
    import torch

    import oracle
    import starkpack_winterfell_b200 as w

    w._lib.init(0)
    for name, field, log_n in (("cfg1_domain_2^23_f128", F128, 23), ("cfg2_domain_2^25_f64", F64, 25)):
        n = 1 << log_n
        ev = fx.synth_f64(0xF2, 1, n)[0] if field == F64 else fx.synth_f128(0xF1, 1, n)[0]
        opts = w.fri.FriOptions(8, 4, 31)
        dev = torch.from_numpy(ev.view(np.int64)).cuda()
        times, roots = [], None
        for rep in range(a.reps + 2):
            ch = HashChannel(field)
            p = w.fri.FriProver(field, B3, opts)
            l0 = w._lib.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            p.build_layers(ch, dev)
            e1.record()
            torch.cuda.synchronize()
            if rep >= 2:
                times.append(e0.elapsed_time(e1))
            roots = ch.commitments
            launches = w._lib.launch_count() - l0
        line = dict(case=name, layers=opts.num_fri_layers(n), folding=4, gpu_ms_median=float(np.median(times)), gpu_ms_min=float(min(times)),
                    gpu_launches=int(launches), bytes_first_layer=int(n * (8 if field == F64 else 16)), root0=roots[0].hex())
        if a.cpu:
            ch = HashChannel(field)
            cur = ev
            t0 = time.perf_counter()
            for _ in range(opts.num_fri_layers(n)):
                l = oracle.fri_commit_layer(field, B3, cur, 4)
                ch.commit_fri_layer(l["root"])
                cur = oracle.fri_apply_drp(field, l["transposed"], 4, opts.domain_offset(field), ch.draw_fri_alpha())
            line["cpu_oracle_ms"] = (time.perf_counter() - t0) * 1e3
            line["cpu_threads"] = f"hashing/merkle {oracle.lib().or_threads()}, folding 1"
            line["roots_equal"] = ch.commitments == roots[: len(ch.commitments)]
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
