#!/bin/bash
# round 2, job E (1 GPU): full -m gpu suite after the fold changes, then bench N=1 (C4 + C2 only) and the pre-binning probe
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest_gpu.log 2>&1
echo "suite rc=$?" | tee -a gpurun_out/e_pytest_gpu.log
tail -8 gpurun_out/e_pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --no-ref-cuda > gpurun_out/e_bench_n1.json 2> gpurun_out/e_bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
l = json.loads(open("gpurun_out/e_bench_n1.json").read().strip().splitlines()[-1])
print("N=1 step %.4f collide %.4f fold %.4f e2e %.4f default %.4f | c2 stream %.5f default %.5f fold %.5f e2e %.5f" % (
    l["ms_per_step"], l["roofline"]["kernel_ms_avg"], l["insert"]["ms_per_frame"], l["e2e"]["ms_per_step"], l["default_path"]["ms_per_step"],
    l["c2"]["streaming"]["ms_per_step"], l["c2"]["default"]["ms_per_step"], l["c2"]["insert_ms"], l["c2"]["e2e"]["ms_per_step"]))
PY
timeout 300 python benchmarks/prebin_probe.py > gpurun_out/e_prebin.log 2>&1; tail -1 gpurun_out/e_prebin.log | head -c 900
