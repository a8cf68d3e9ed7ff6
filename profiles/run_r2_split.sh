#!/bin/bash
# round 2: the split plan (transforms above the two-pass limit, and preferred for the LDE of 2^23 / 2^24-row traces) on one
# GPU: parity tests, then a 2^24-row x 32-column shard (cfg5's per-GPU share at 8 GPUs) and a 2^23-row x 64-column one with
# the preference on (default, 22) and off (24)
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_split.py tests/test_gpu_sharded.py -x -q > gpurun_out/pytest_split.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_split.log
tail -12 gpurun_out/pytest_split.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1"
rm -f gpurun_out/split_1gpu.jsonl
for lr in 24 23; do for sp in 24 22; do
  timeout 600 $TR --master-port 296$lr profiles/bench_sharded.py --log-rows $lr --cols $((32 * (25 - lr))) --reps 2 --forms 1:1 --split-pref $sp >> gpurun_out/split_1gpu.jsonl 2>> gpurun_out/split_1gpu.err
done; done
cat gpurun_out/split_1gpu.jsonl; tail -3 gpurun_out/split_1gpu.err
