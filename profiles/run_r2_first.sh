#!/bin/bash
# round 2, first GPU call: integer-pipe micro-benchmark, the new sharded tests, the headline bench on one GPU
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.csv
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt
cd profiles
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o /tmp/int_peak int_peak.cu && /tmp/int_peak > ../gpurun_out/int_peak.jsonl 2> ../gpurun_out/int_peak.err
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -DWG_MUL_I128 -o /tmp/int_peak_i128 int_peak.cu && /tmp/int_peak_i128 > ../gpurun_out/int_peak_i128.jsonl 2>> ../gpurun_out/int_peak.err
cd ..
timeout 600 python profiles/ab_variants.py base i128 dit dit_i128 > gpurun_out/ab_variants.jsonl 2> gpurun_out/ab_variants.err
cat gpurun_out/ab_variants.jsonl
timeout 1500 python -m pytest tests/test_gpu_parity.py::test_device_lazy_goldilocks_ops tests/test_gpu_parity.py::test_device_field_ops_f64 tests/test_gpu_sharded.py tests/test_gpu_parallel.py tests/test_gpu_fullsize.py -x -q > gpurun_out/pytest_sharded.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_sharded.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
echo "bench rc=$?" >> gpurun_out/bench_1gpu.err
tail -5 gpurun_out/pytest_sharded.log; tail -3 gpurun_out/bench_1gpu.err; head -c 600 gpurun_out/bench_1gpu.json
