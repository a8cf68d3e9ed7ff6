"""Profiling driver: a reduced-width copy of the bench workload (2^22 rows x 16 f64 columns, blowup 8, Blake3_256)
so that every kernel of the commit runs with its production tile shape but an ncu replay stays short.
usage: python profiles/prof_commit.py [commits]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import starkpack_winterfell_b200 as w
from starkpack_winterfell_b200.device import TraceCommitter

cols = int(os.environ.get("PROF_COLS", "16"))
log_rows = int(os.environ.get("PROF_LOG_ROWS", "22"))
n = 1 << log_rows
tc = TraceCommitter(w.F64, w.BLAKE3_256, n, 8, [cols], device="cuda:0")
tc.cols.copy_(bench.synth_trace_device(torch, cols, n, bench.SEED, tc.device))
torch.cuda.synchronize()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    tc.commit()
torch.cuda.synchronize()
print("root", tc.root().hex())
