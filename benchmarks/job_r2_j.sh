#!/bin/bash
# collision lanes: their tests (with output), the C++ example's timing lines, then the whole -m gpu suite
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lanes.py -m gpu -x -q -s > gpurun_out/j_lanes.log 2>&1
echo "lanes rc=$?" | tee -a gpurun_out/j_lanes.log
tail -25 gpurun_out/j_lanes.log
for k in 1 2 3; do timeout 120 tests/cpp/lanes_example; done > gpurun_out/j_lanes_example.log 2>&1
echo "example rc=$?"; cat gpurun_out/j_lanes_example.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/j_pytest_gpu.log
tail -8 gpurun_out/j_pytest_gpu.log
