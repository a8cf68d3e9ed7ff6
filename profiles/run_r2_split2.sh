#!/bin/bash
# split plan on narrow column chunks (what a sharded commit extends at a time): 2^24 rows x 8 / 16 columns, preference on / off
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_split.py tests/test_gpu_sharded.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_split2.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_split2.log; tail -4 gpurun_out/pytest_split2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1"
rm -f gpurun_out/split_narrow_1gpu.jsonl
for c in 8 16; do for sp in 24 22; do
  timeout 600 $TR --master-port 2970$((c/8)) profiles/bench_sharded.py --log-rows 24 --cols $c --reps 2 --forms 1:1 --split-pref $sp >> gpurun_out/split_narrow_1gpu.jsonl 2>> gpurun_out/split_narrow_1gpu.err
done; done
grep n_gpus gpurun_out/split_narrow_1gpu.jsonl; tail -3 gpurun_out/split_narrow_1gpu.err
