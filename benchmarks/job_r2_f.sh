#!/bin/bash
# round 2, job F (1 GPU): sensor/parity tests + the fold on a 1/8 z-slab (what a rank does at N = 8) and on the whole grid
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q > gpurun_out/f_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/f_pytest.log
for DZ in 256 2048 f; do
  EXTRA=""; [ $DZ = f ] && EXTRA="--z0 4096" && DZARG=256 || DZARG=$DZ
  timeout 300 python benchmarks/prebin_probe.py --dz $DZARG $EXTRA --out gpurun_out/f_prebin_$DZ.json > gpurun_out/f_prebin_$DZ.log 2>&1
  python - <<PY
import json
r = json.load(open("gpurun_out/f_prebin_$DZ.json"))
print("dz=$DZ fold steady %.4f ms (min %.4f, back to back %.4f) first frame %.4f presorted %.4f identical %s" % (r["A_unsorted_steady_ms"]["median"], r["A_unsorted_steady_ms"]["min"], r["A_unsorted_back_to_back_ms"], r["A_unsorted_first_frame_ms"], r["B_presorted_steady_ms"]["median"], r["maps_identical"]))
PY
done
