# round-1 closing run: all GPU tests, smoke, the default bench, the TMA A/B probe (last: its bounded waits trap on failure)
set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python benchmarks/tma_probe.py 40 > gpurun_out/tma_probe.log 2>&1; echo "probe rc=$?"
tail -5 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; tail -20 gpurun_out/tma_probe.log
