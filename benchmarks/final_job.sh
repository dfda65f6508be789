# round-1 closing run: all GPU tests, smoke, the default bench, the TMA A/B probe, ncu of the TMA kernel
set -x
( time timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python benchmarks/tma_probe.py 40 > gpurun_out/tma_probe.log 2>&1; echo "probe rc=$?"
timeout 600 python benchmarks/tma_probe.py 4 bitprob > gpurun_out/tma_probe_short.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_collide_tma -s 6 -c 1 -f -o gpurun_out/r1_collide_tma python benchmarks/tma_probe.py 4 bitprob > gpurun_out/ncu_tma.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; tail -20 gpurun_out/tma_probe.log; ls -la gpurun_out
