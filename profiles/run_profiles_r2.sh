#!/bin/bash
# round 2 profiling pass (one GPU): plain bench first (must exit 0), then the ncu launch list of the same command, one
# `--set full` capture of the commit kernels (reduced width: production tile shapes, short replay), and a per-launch
# traffic + executed-instruction pass at the FULL bench size.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-configs > gpurun_out/r2_bench_plain.json 2> gpurun_out/r2_bench_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/r2_bench_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ntt_pass|hash_rows|merkle|twiddles" -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-configs > gpurun_out/ncu_launch.log 2>&1
timeout 200 python profiles/prof_commit.py 1 > gpurun_out/prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ntt_pass|hash_rows|merkle_level" -c 7 -o gpurun_out/r2_full python profiles/prof_commit.py 1 > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r2_full.ncu-rep --page raw --csv > gpurun_out/r2_full_raw.csv 2> gpurun_out/ncu_export.err
python profiles/summarize_ncu.py gpurun_out/r2_full_raw.csv > gpurun_out/r2_ncu_full_summary.txt 2>> gpurun_out/ncu_export.err
timeout 200 python profiles/phase_times.py 1 > gpurun_out/phase_plain.log 2>&1 &&
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum --clock-control none -k regex:"ntt_pass|hash_rows|merkle_level" -c 8 --csv --log-file gpurun_out/r2_traffic.csv python profiles/phase_times.py 1 > gpurun_out/ncu_traffic.log 2>&1
rm -f gpurun_out/r2_full.ncu-rep.tmp
ls -la gpurun_out/ | tail -20
head -40 gpurun_out/r2_ncu_full_summary.txt
