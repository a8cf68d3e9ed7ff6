"""Tuning driver: device-resident commit of the bench workload (2^22 x 100 f64, blowup 8, Blake3_256), prints the
per-phase CUDA-event times and the root. Environment knobs of the library (WG_*) apply.
usage: python profiles/phase_times.py [commits] [log_rows] [cols]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import starkpack_winterfell_b200 as w
from starkpack_winterfell_b200 import _lib
from starkpack_winterfell_b200.device import TraceCommitter

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
log_rows = int(sys.argv[2]) if len(sys.argv) > 2 else 22
cols = int(sys.argv[3]) if len(sys.argv) > 3 else 100
n = 1 << log_rows
tc = TraceCommitter(w.F64, w.BLAKE3_256, n, 8, [cols], device="cuda:0")
tc.cols.copy_(bench.synth_trace_device(torch, cols, n, bench.SEED, tc.device))
torch.cuda.synchronize()
tc.commit()
torch.cuda.synchronize()
_lib.profile_enable(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    tc.commit()
e1.record()
torch.cuda.synchronize()
ph = _lib.profile_read()
_lib.profile_enable(False)
knobs = {k: v for k, v in os.environ.items() if k.startswith("WG_")}
print(knobs, "ms/commit %.2f" % (e0.elapsed_time(e1) / reps), {k: round(v[0] / reps, 2) for k, v in ph.items()}, tc.root().hex()[:16])
