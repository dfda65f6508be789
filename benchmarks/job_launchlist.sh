# launch list of the bench command under ncu (durations cold-cache and serialised: shares, not absolutes)
timeout 40 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 && \
timeout 50 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1_launches_bench_summary.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
grep -c k_collide_summary2 gpurun_out/r1_launches_bench_summary.csv
