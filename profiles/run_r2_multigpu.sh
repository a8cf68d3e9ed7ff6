#!/bin/bash
# round 2 multi-GPU evidence: usage  bash profiles/run_r2_multigpu.sh <N>
# (1) sharded commit over real NCCL / P2P against the oracle on small shapes, (2) bench.py --gpus N (strong scaling of the
# headline trace through wg_sharded_commit_dev + cfg4 / cfg5 sub-benchmarks), (3) cfg5 full-size property check.
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_${N}.csv
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29611 tests/run_distributed_gpu.py > gpurun_out/dist_small_${N}gpu.log 2>&1
echo "rc=$?" >> gpurun_out/dist_small_${N}gpu.log; tail -3 gpurun_out/dist_small_${N}gpu.log
timeout 1500 $TR --master-port 29612 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
echo "rc=$?" >> gpurun_out/bench_${N}gpu.err; tail -5 gpurun_out/bench_${N}gpu.err; head -c 400 gpurun_out/bench_${N}gpu.json
timeout 1200 $TR --master-port 29613 tests/run_distributed_gpu.py --cfg5 > gpurun_out/cfg5_${N}gpu.jsonl 2> gpurun_out/cfg5_${N}gpu.err
echo "rc=$?" >> gpurun_out/cfg5_${N}gpu.err; tail -3 gpurun_out/cfg5_${N}gpu.err; cat gpurun_out/cfg5_${N}gpu.jsonl
[ "$N" -ge 8 ] && exit 0   # (the reference arm does not depend on N; not repeated on the 8-GPU box)
timeout 600 $TR --master-port 29614 bench.py --impl reference --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_ref_${N}gpu.json 2> gpurun_out/bench_ref_${N}gpu.err
echo "rc=$?" >> gpurun_out/bench_ref_${N}gpu.err; head -c 300 gpurun_out/bench_ref_${N}gpu.json
