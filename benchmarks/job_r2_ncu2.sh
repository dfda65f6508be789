#!/bin/bash
# round 2, final ncu job (1 GPU): launch list of the C4 step + --set full of the final fold kernel (whole grid, 1/8 slab)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
BENCH="python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $BENCH > gpurun_out/n2_bench_c4.json 2> gpurun_out/n2_bench_c4.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c4.csv $BENCH > gpurun_out/n2_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_update_logodds_frames --launch-skip 6 -c 1 -f -o gpurun_out/r2_fold_n1 $BENCH > gpurun_out/n2_ncu_fold.log 2>&1; echo "ncu fold rc=$?"
ncu -i gpurun_out/r2_fold_n1.ncu-rep --page raw --csv > gpurun_out/r2_fold_n1_ncu_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_update_logodds_frames --launch-skip 8 -c 1 -f -o gpurun_out/r2_fold_slab python benchmarks/prebin_probe.py --dz 256 --out gpurun_out/n2_prebin.json > gpurun_out/n2_ncu_slab.log 2>&1; echo "ncu slab rc=$?"
ncu -i gpurun_out/r2_fold_slab.ncu-rep --page raw --csv > gpurun_out/r2_fold_slab_ncu_raw.csv 2>/dev/null
ncu -i gpurun_out/r2_fold_slab.ncu-rep --page source --csv > gpurun_out/r2_fold_slab_ncu_source.csv 2>/dev/null
ls -la gpurun_out/r2_fold* gpurun_out/r2_launches_bench_c4.csv
