#!/bin/bash
# ncu --set full of the multi-frame fold on a 1/8 z-slab (a rank's work at N = 8), steady state
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_update_logodds_frames --launch-skip 8 -c 1 -f -o gpurun_out/r2_fold_slab python benchmarks/prebin_probe.py --dz 256 --out gpurun_out/g_prebin.json > gpurun_out/g_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r2_fold_slab.ncu-rep --page raw --csv > gpurun_out/r2_fold_slab_ncu_raw.csv 2>/dev/null
ncu -i gpurun_out/r2_fold_slab.ncu-rep --page source --csv > gpurun_out/r2_fold_slab_ncu_source.csv 2>/dev/null
ls -la gpurun_out/r2_fold_slab*
