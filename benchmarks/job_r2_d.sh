#!/bin/bash
# round 2, job D (1 GPU): the no-alloc test alone in a fresh process, then the whole -m gpu suite (it runs last there)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zzz_no_alloc.py -m gpu -x -q > gpurun_out/d_noalloc_alone.log 2>&1
echo "alone rc=$?" | tee -a gpurun_out/d_noalloc_alone.log
tail -15 gpurun_out/d_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest_gpu.log 2>&1
echo "suite rc=$?" | tee -a gpurun_out/d_pytest_gpu.log
tail -15 gpurun_out/d_pytest_gpu.log
