#!/bin/bash
# multi_gpu_check only, at world $1
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N --master-port 29511 tests/multi_gpu_check.py > gpurun_out/m_check_n$N.log 2>&1
echo "multi_gpu_check rc=$?" | tee -a gpurun_out/m_check_n$N.log
tail -4 gpurun_out/m_check_n$N.log
