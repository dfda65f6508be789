#!/bin/bash
# per-meaning list x map as ONE launch: the whole -m gpu suite, then the C5 bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/p_pytest_gpu.log
tail -3 gpurun_out/p_pytest_gpu.log
timeout 300 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > gpurun_out/p_bench_c5.json 2> gpurun_out/p_bench_c5.err; echo "bench rc=$?"
python - <<PY
import json
l = json.loads(open("gpurun_out/p_bench_c5.json").read().strip().splitlines()[-1])
for t in ("streaming", "default"):
    c = l["c5"][t]; print(t, c["ms_per_step"], c["count_only_ms"], c["local_kernel_ms"], c["parity"]["total_equal"], c["parity"]["per_meaning_equal"])
PY
