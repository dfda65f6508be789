# round-1 scaling run on one 8-GPU box: the driver's own launch line for N=2,4,8, then the peer-memory checks
set -x
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 100 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 100 --warmup 3 --diag > /dev/null 2> gpurun_out/diag_n$n.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 tests/multi_gpu_check.py > gpurun_out/mgc8.log 2>&1; echo "mgc8 rc=$?" >> gpurun_out/mgc8.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 tests/multi_gpu_check.py > gpurun_out/mgc2.log 2>&1; echo "mgc2 rc=$?" >> gpurun_out/mgc2.log
python bench.py --steps 100 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_samebox.json 2>/dev/null
tail -3 gpurun_out/mgc8.log gpurun_out/mgc2.log; cat gpurun_out/bench_n*.json | cut -c1-220
