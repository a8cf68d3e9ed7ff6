#!/bin/bash
# final build on 2 GPUs: every exchange form against the oracle, then the driver's N = 2 bench command
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29721 tests/run_distributed_gpu.py > gpurun_out/dist_small_2gpu_f.log 2>&1
echo "rc=$?" >> gpurun_out/dist_small_2gpu_f.log; tail -3 gpurun_out/dist_small_2gpu_f.log
timeout 900 $TR --master-port 29722 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu_f.json 2> gpurun_out/bench_2gpu_f.err
echo "rc=$?" >> gpurun_out/bench_2gpu_f.err; tail -4 gpurun_out/bench_2gpu_f.err; head -c 400 gpurun_out/bench_2gpu_f.json
