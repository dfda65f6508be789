#!/bin/bash
# the whole -m gpu suite
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c_pytest_gpu.log
tail -15 gpurun_out/c_pytest_gpu.log
