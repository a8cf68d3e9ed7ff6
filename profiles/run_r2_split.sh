#!/bin/bash
# round 2: the split plan (transforms above the two-pass limit) on one GPU: parity tests, then a 2^24-row x 32-column shard
# (cfg5's per-GPU share at 8 GPUs) with the default 2^12-point passes against radix-4 + 2^11-point passes (--split-log 22)
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_split.py -x -q > gpurun_out/pytest_split.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_split.log
tail -12 gpurun_out/pytest_split.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1"
for sl in 24 22 23; do
  timeout 600 $TR --master-port 2963$((sl-20)) profiles/bench_sharded.py --log-rows 24 --cols 32 --reps 2 --forms 1:1 --split-log $sl >> gpurun_out/split_1gpu.jsonl 2>> gpurun_out/split_1gpu.err
done
cat gpurun_out/split_1gpu.jsonl; tail -3 gpurun_out/split_1gpu.err
