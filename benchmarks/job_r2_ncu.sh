#!/bin/bash
# round 2, ncu job (1 GPU): launch list of the bench's C4 step, --set full of the dominant kernels, pre-binning probe
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python benchmarks/prebin_probe.py > gpurun_out/n_prebin.log 2>&1; echo "prebin rc=$?"; tail -2 gpurun_out/n_prebin.log | head -c 1500
BENCH="python bench.py --config c4 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $BENCH > gpurun_out/n_bench_c4.json 2> gpurun_out/n_bench_c4.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c4.csv $BENCH > gpurun_out/n_ncu_list.log 2>&1; echo "ncu list rc=$?"
for K in k_collide_prob_prob k_update_logodds_frames k_collide_summary2; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K --launch-skip 6 -c 1 -f -o gpurun_out/r2_$K $BENCH > gpurun_out/n_ncu_$K.log 2>&1; echo "ncu $K rc=$?"
  ncu -i gpurun_out/r2_$K.ncu-rep --page raw --csv > gpurun_out/r2_${K}_ncu_raw.csv 2>/dev/null
done
BENCH5="python bench.py --config c5 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_collide_list_map --launch-skip 8 -c 2 -f -o gpurun_out/r2_k_collide_list_map $BENCH5 > gpurun_out/n_ncu_list_map.log 2>&1; echo "ncu list_map rc=$?"
ncu -i gpurun_out/r2_k_collide_list_map.ncu-rep --page raw --csv > gpurun_out/r2_k_collide_list_map_ncu_raw.csv 2>/dev/null
ls -la gpurun_out | tail -20
