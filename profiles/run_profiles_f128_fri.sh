# ncu evidence for the f128 path (cfg1 shape) and the FRI kernels; run on one B200:  gpurun -- bash profiles/run_profiles_f128_fri.sh
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
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"ntt_pass|hash_rows" -s 5 -c 5 -o gpurun_out/r1_f128_full python /tmp/prof_f128.py > gpurun_out/ncu_f128.log 2>&1
ncu -i gpurun_out/r1_f128_full.ncu-rep --page raw --csv > gpurun_out/r1_f128_raw.csv 2>/dev/null
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"fri_|hash_rows|merkle|ntt_pass|scale_powers" -c 120 --csv --log-file gpurun_out/r1_fri_launches.csv python profiles/bench_fri.py --reps 1 --cpu 0 > gpurun_out/ncu_fri.log 2>&1
rm -f gpurun_out/r1_f128_full.ncu-rep
ls -la gpurun_out | tail -8
