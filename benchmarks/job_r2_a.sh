#!/bin/bash
# round 2, job A (1 GPU): new parity tests, then the first full bench line + the reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/a_gpu.txt; nproc >> gpurun_out/a_gpu.txt; free -g >> gpurun_out/a_gpu.txt
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_large.py -m gpu -x -q > gpurun_out/a_pytest_new.log 2>&1
echo "pytest new rc=$?" >> gpurun_out/a_pytest_new.log
tail -5 gpurun_out/a_pytest_new.log
timeout 600 python bench.py > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err
echo "bench rc=$?"; tail -c 1500 gpurun_out/a_bench_n1.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err
echo "ref rc=$?"
head -c 3000 gpurun_out/a_bench_n1.json
