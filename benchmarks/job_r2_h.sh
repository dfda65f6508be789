#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for G in 0 32 128; do for DZ in 256 2048; do
  timeout 300 python benchmarks/prebin_probe.py --dz $DZ --gran $G --iters 20 --out gpurun_out/h_$G_$DZ.json > gpurun_out/h.log 2>&1
  python - <<PY
import json
r = json.load(open("gpurun_out/h_$G_$DZ.json"))
print("gran=$G (default %s) dz=$DZ fold back-to-back %.4f ms, single %.4f, first frame %.4f" % (r["l2_fetch_granularity_default"], r["A_unsorted_back_to_back_ms"], r["A_unsorted_steady_ms"]["median"], r["A_unsorted_first_frame_ms"]))
PY
done; done
