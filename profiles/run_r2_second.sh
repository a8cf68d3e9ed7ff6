#!/bin/bash
# round 2: full GPU test suite, headline bench on one GPU, integer-pipe micro-benchmark
set -x
mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
echo "bench rc=$?" >> gpurun_out/bench_1gpu.err
tail -3 gpurun_out/bench_1gpu.err
cd profiles
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o /tmp/int_peak int_peak.cu && /tmp/int_peak > ../gpurun_out/int_peak.jsonl 2> ../gpurun_out/int_peak.err
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -DWG_MUL_I128 -o /tmp/int_peak_i128 int_peak.cu && /tmp/int_peak_i128 > ../gpurun_out/int_peak_i128.jsonl 2>> ../gpurun_out/int_peak.err
tail -2 ../gpurun_out/int_peak_i128.jsonl
