#!/bin/bash
# round 2, final one-GPU pass on the final build: full GPU suite, the no-allocation test alone, the default bench line
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_final.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu_final.log
tail -6 gpurun_out/pytest_gpu_final.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone_final.log 2>&1; tail -2 gpurun_out/noalloc_alone_final.log
timeout 900 python bench.py > gpurun_out/bench_1gpu_final.json 2> gpurun_out/bench_1gpu_final.err
echo "bench rc=$?" >> gpurun_out/bench_1gpu_final.err; tail -3 gpurun_out/bench_1gpu_final.err; head -c 600 gpurun_out/bench_1gpu_final.json
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_1gpu_final.json 2> gpurun_out/bench_ref_1gpu_final.err
echo "ref rc=$?" >> gpurun_out/bench_ref_1gpu_final.err; head -c 600 gpurun_out/bench_ref_1gpu_final.json
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/smoke_final.log 2>&1; tail -2 gpurun_out/smoke_final.log
