#!/bin/bash
# round 2, third one-GPU pass: full GPU suite on the final build, f128 out-of-line-product A/B, then the profiling pass
# (plain bench -> ncu launch list -> --set full capture -> traffic / executed-instruction pass).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_c.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_c.log
tail -6 gpurun_out/pytest_gpu_c.log
timeout 600 python profiles/ab_f128.py cur f128call > gpurun_out/ab_f128.jsonl 2> gpurun_out/ab_f128.err
cat gpurun_out/ab_f128.jsonl
bash profiles/run_profiles_r2.sh
