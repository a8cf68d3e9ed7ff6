#!/bin/bash
# round 2: the copy-engine exchange (WG_EXCHANGE_P2P_COPY) against the NCCL forms.  usage: bash profiles/run_r2_forms.sh <N> [tag]
# (1) small shapes against the oracle on N GPUs with every form (tests/run_distributed_gpu.py), (2) the headline trace with
# pipelined NCCL (2:K), copy-engine pushes (4:K) and the auto policy, (3) at N >= 4 the same for the cfg5 shape (2^24 x 256).
N=${1:-2}
TAG=${2:-b}
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ -z "$SKIP_SMALL" ]; then
timeout 900 $TR --master-port 29621 tests/run_distributed_gpu.py > gpurun_out/dist_small_${N}gpu_${TAG}.log 2>&1
echo "rc=$?" >> gpurun_out/dist_small_${N}gpu_${TAG}.log; tail -3 gpurun_out/dist_small_${N}gpu_${TAG}.log
fi
timeout 600 $TR --master-port 29622 profiles/bench_sharded.py --forms ${FORMS:-2:2,2:4,4:1,4:2,4:4,0:0} > gpurun_out/forms_${N}gpu_${TAG}.jsonl 2> gpurun_out/forms_${N}gpu_${TAG}.err
echo "rc=$?" >> gpurun_out/forms_${N}gpu_${TAG}.err; tail -2 gpurun_out/forms_${N}gpu_${TAG}.err; cat gpurun_out/forms_${N}gpu_${TAG}.jsonl
if [ "$N" -ge 4 ]; then
  timeout 900 $TR --master-port 29623 profiles/bench_sharded.py --log-rows 24 --cols 256 --reps 2 --forms ${FORMS5:-2:4,4:2,4:4,0:0} > gpurun_out/forms_cfg5_${N}gpu_${TAG}.jsonl 2> gpurun_out/forms_cfg5_${N}gpu_${TAG}.err
  # the same with the split plan's preference off (two 2^12-point passes)
  timeout 900 $TR --master-port 29624 profiles/bench_sharded.py --log-rows 24 --cols 256 --reps 2 --forms 4:4 --split-pref 24 >> gpurun_out/forms_cfg5_${N}gpu_${TAG}.jsonl 2>> gpurun_out/forms_cfg5_${N}gpu_${TAG}.err
  echo "rc=$?" >> gpurun_out/forms_cfg5_${N}gpu_${TAG}.err; tail -2 gpurun_out/forms_cfg5_${N}gpu_${TAG}.err; cat gpurun_out/forms_cfg5_${N}gpu_${TAG}.jsonl
fi
exit 0
