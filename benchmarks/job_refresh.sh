# First GPU call of a new round (N=1, about 3 minutes of box time): everything under profiles/ that predates the last
# kernel changes of round 1 or has never run on a GPU.
#   gpurun --timeout 600 -- 'bash benchmarks/job_refresh.sh'
# 1. the whole GPU suite, the late file (tests/test_zz_cpp_facade_late.py) included -- its first run
( time timeout 600 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
# 2. both bench arms
timeout 300 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
# 3. the five BASELINE configs (profiles/r1_configs.json predates PDL and the one-atomic finish)
timeout 400 python benchmarks/run_configs.py --out gpurun_out/configs.json > gpurun_out/configs.log 2>&1; echo "configs rc=$?"
python - <<'PY'
import json
b = json.load(open("gpurun_out/bench_n1.json")); r = b["roofline"]
print("value %.2f T voxel/s, step %.1f us, e2e %.1f us, collide %.2f us (frac %.2f), insert %.2f us" % (
    b["value"] / 1e12, b["ms_per_step"] * 1e3, b["e2e"]["ms_per_step"] * 1e3, r["kernel_ms_avg"] * 1e3, r["frac"],
    b["insert"]["ms_per_frame"] * 1e3))
PY
