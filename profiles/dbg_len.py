import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import oracle
from tests import fixtures as fx
from starkpack_winterfell_b200.device import TraceCommitter
log_n = int(sys.argv[1]); cols = int(sys.argv[2]); B = 8
n = 1 << log_n
tr = fx.synth_f64(0xA000 + log_n, cols, n)
tc = TraceCommitter(0, 0, n, B, [cols], device="cuda:0")
tc.cols.copy_(torch.from_numpy(np.ascontiguousarray(tr).view(np.int64).reshape(tc.cols.shape)).to(tc.device))
tc.commit(); torch.cuda.synchronize()
oracle.build()
want = oracle.commit_trace(0, 0, [tr], B, want=("lde",))
got = tc.ldes[0].cpu().numpy().view(np.uint64).reshape(want["lde"][0].shape)
bad = np.argwhere(got != want["lde"][0])
print({k: v for k, v in os.environ.items() if k.startswith("WG_")}, "log_n", log_n, "mismatches", len(bad), "of", got.size)
if len(bad):
    rows = np.unique(bad[:, 0]); colsb = np.unique(bad[:, 1])
    print(" first rows", rows[:10], "cols", colsb, "rows%8", np.unique(rows % 8), "row>>3 min/max", (rows >> 3).min(), (rows >> 3).max())
    j = rows >> 3
    print(" j % 2048 uniq count", len(np.unique(j % 2048)), " j//2048 uniq", np.unique(j // 2048)[:20], " j%4096 cnt", len(np.unique(j % 4096)), "j//4096", np.unique(j//4096)[:10])
