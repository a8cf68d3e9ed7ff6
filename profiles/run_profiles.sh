mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || echo "bench failed"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ntt_pass|hash_rows|merkle|twiddles|fri_|interleave" -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"ntt_pass|hash_rows|merkle_level" -c 7 -o gpurun_out/r1_full python profiles/prof_commit.py 1 > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r1_full.ncu-rep --page raw --csv > gpurun_out/r1_full_raw.csv 2> gpurun_out/ncu_export.err
ls -la gpurun_out/
timeout 400 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"ntt_pass|hash_rows|merkle_level" -c 8 --csv --log-file gpurun_out/traffic_r1.csv python profiles/phase_times.py 1 > gpurun_out/ncu_traffic.log 2>&1
rm -f gpurun_out/r1_full.ncu-rep.tmp
du -sh gpurun_out/*
head -c 400 gpurun_out/bench_final.json
