#!/bin/bash
# round 2, last ncu job (1 GPU): --set full of k_collide_list_map in per-meaning mode after its tail gained the in-kernel
# 257-word all-reduce (main loop unchanged), then the whole -m gpu suite and smoke()
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
BENCH5="python bench.py --config c5 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $BENCH5 > gpurun_out/n3_bench_c5.json 2> gpurun_out/n3_bench_c5.err; echo "bench c5 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_collide_list_map --launch-skip 8 -c 2 -f -o gpurun_out/r2_k_collide_list_map $BENCH5 > gpurun_out/n3_ncu_list_map.log 2>&1; echo "ncu list_map rc=$?"
ncu -i gpurun_out/r2_k_collide_list_map.ncu-rep --page raw --csv > gpurun_out/r2_k_collide_list_map_ncu_raw.csv 2>/dev/null
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/n3_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/n3_pytest_gpu.log
tail -4 gpurun_out/n3_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/n3_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/n3_smoke.log
