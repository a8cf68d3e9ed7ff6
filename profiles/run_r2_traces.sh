#!/bin/bash
# round 2: several narrow traces extended together (lde_traces_dev): parity tests of every multi-trace entry, then the f128
# configs (cfg1, cfg3, cfg4 = 8 traces of 2 columns) timed through the device entry
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_containers.py tests/test_gpu_parity.py tests/test_gpu_resident.py tests/test_gpu_fullsize.py tests/test_gpu_split.py -x -q > gpurun_out/pytest_traces.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_traces.log; tail -5 gpurun_out/pytest_traces.log
timeout 600 python profiles/ab_f128.py main > gpurun_out/f128_configs_main.jsonl 2> gpurun_out/f128_configs_main.err; cat gpurun_out/f128_configs_main.jsonl
