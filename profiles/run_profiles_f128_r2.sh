#!/bin/bash
# round 2 ncu evidence for the f128 path (cfg1 shape: fib 2^20 x 2): plain run first, then one --set full capture of the
# commit kernels, summarised by summarize_ncu.py
set -x
mkdir -p gpurun_out
cat > /tmp/prof_f128.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import starkpack_winterfell_b200 as w
from starkpack_winterfell_b200.device import TraceCommitter
from tests import fixtures as fx
n = 1 << 20
tc = TraceCommitter(w.F128, w.BLAKE3_256, n, 8, [2], device="cuda:0")
tr = fx.fib_trace_f128(n)
tc.cols.copy_(torch.from_numpy(np.ascontiguousarray(tr).view(np.int64).reshape(2, n, 2)).cuda())
for _ in range(2):
    tc.commit()
torch.cuda.synchronize()
print("root", tc.root().hex())
PY
timeout 200 python /tmp/prof_f128.py > gpurun_out/prof_f128_plain.log 2>&1 || { echo plain run failed; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"ntt_pass|hash_rows" -s 5 -c 5 -o gpurun_out/r2_f128_full python /tmp/prof_f128.py > gpurun_out/ncu_f128.log 2>&1
ncu -i gpurun_out/r2_f128_full.ncu-rep --page raw --csv > gpurun_out/r2_f128_raw.csv 2>/dev/null
python profiles/summarize_ncu.py gpurun_out/r2_f128_raw.csv > gpurun_out/r2_ncu_f128_summary.txt
rm -f gpurun_out/r2_f128_full.ncu-rep
head -30 gpurun_out/r2_ncu_f128_summary.txt
