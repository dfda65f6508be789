# round-1 final captures: plain runs first (their numbers are the bench values), then the launch list and full captures
set -x
python bench.py --steps 100 --warmup 3 > gpurun_out/b1.json 2> gpurun_out/b1.err || exit 1
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/bref.json 2> gpurun_out/bref.err
python benchmarks/run_configs.py --out gpurun_out/r1_configs.json > gpurun_out/configs.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches_bench_summary.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_collide_summary2 -s 6 -c 2 -f -o gpurun_out/r1_collide_summary2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_collide_s2.log 2>&1
python benchmarks/run_configs.py --configs c4s,c2r --out gpurun_out/tmp_cfg.json > gpurun_out/ncu_plain_cfg.log 2>&1 || exit 2
ncu --set full --clock-control none --import-source on -k regex:k_collide_summary2 -s 8 -c 1 -f -o gpurun_out/r1_collide_summary2_c4s python benchmarks/run_configs.py --configs c4s --out gpurun_out/tmp_cfg.json > gpurun_out/ncu_c4s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_raycast_free -s 3 -c 2 -f -o gpurun_out/r1_raycast python benchmarks/run_configs.py --configs c2r --out gpurun_out/tmp_cfg.json > gpurun_out/ncu_c2r.log 2>&1
ls -la gpurun_out/
