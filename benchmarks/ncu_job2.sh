# round-1 captures after the summary redesign: plain run first, then the launch list, then one full capture
set -x
python bench.py --steps 100 --warmup 3 > gpurun_out/b1.json 2> gpurun_out/b1.err || exit 1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches_bench_summary.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_collide_summary2 -s 6 -c 2 -f -o gpurun_out/r1_collide_summary2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_collide_s2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_update_logodds -s 6 -c 2 -f -o gpurun_out/r1_update_logodds python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_upd.log 2>&1
ls -la gpurun_out/
