#!/bin/bash
# the driver's N = 4 scaling step with the final build: bench.py --gpus 4 (headline sharded + cfg4 + cfg5 sub-benchmarks)
# and the cfg5 full-size property check
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 1200 $TR --master-port 29712 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/bench_4gpu_f.json 2> gpurun_out/bench_4gpu_f.err
echo "rc=$?" >> gpurun_out/bench_4gpu_f.err; tail -4 gpurun_out/bench_4gpu_f.err; head -c 300 gpurun_out/bench_4gpu_f.json
timeout 900 $TR --master-port 29713 tests/run_distributed_gpu.py --cfg5 > gpurun_out/cfg5_4gpu_f.jsonl 2> gpurun_out/cfg5_4gpu_f.err
echo "rc=$?" >> gpurun_out/cfg5_4gpu_f.err; tail -3 gpurun_out/cfg5_4gpu_f.err; cat gpurun_out/cfg5_4gpu_f.jsonl
