#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_coarse.py tests/test_gpu_summary.py tests/test_gpu_parity.py tests/test_gpu_round2.py tests/test_gpu_sweptvolume.py -m gpu -x -q > gpurun_out/i_pytest.log 2>&1
echo "pytest rc=$?"; tail -12 gpurun_out/i_pytest.log
timeout 300 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > gpurun_out/i_bench_c2.json 2> gpurun_out/i_bench_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/i_bench_c2.err
python - <<PY
import json
l = json.loads(open("gpurun_out/i_bench_c2.json").read().strip().splitlines()[-1])
c = l["c2"]
print("c2 default step %.5f ms collide %.5f | streaming step %.5f collide %.5f | fold %.5f | e2e %.5f | counts %s %s" % (c["default"]["ms_per_step"], c["default"]["collide_kernel_ms"], c["streaming"]["ms_per_step"], c["streaming"]["collide_kernel_ms"], c["insert_ms"], c["e2e"]["ms_per_step"], c["default"]["count"], c["streaming"]["count"]))
PY
