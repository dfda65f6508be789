# round-1 scaling run on one 8-GPU box: peer-memory checks first, then the driver's own launch line for N=2,4,8
set -x
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29701 tests/multi_gpu_check.py > gpurun_out/mgc8.log 2>&1; rc=$?; echo "mgc8 rc=$rc" >> gpurun_out/mgc8.log
tail -n 3 gpurun_out/mgc8.log
[ $rc -ne 0 ] && exit 1
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 100 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err || tail -n 5 gpurun_out/bench_n$n.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 bench.py --gpus 8 --steps 100 --warmup 3 --diag > /dev/null 2> gpurun_out/diag_n8.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29609 bench.py --gpus 8 --steps 100 --warmup 3 --fold multi > gpurun_out/bench_n8_multi.json 2> gpurun_out/bench_n8_multi.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29610 bench.py --gpus 4 --steps 100 --warmup 3 --fold multi > gpurun_out/bench_n4_multi.json 2> gpurun_out/bench_n4_multi.err
python bench.py --steps 100 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_samebox.json 2>/dev/null
grep -o "diag rank 0: {[^}]*}" gpurun_out/diag_n8.err
cat gpurun_out/bench_n1_samebox.json gpurun_out/bench_n2.json gpurun_out/bench_n4.json gpurun_out/bench_n8.json gpurun_out/bench_n4_multi.json gpurun_out/bench_n8_multi.json | python -c "
import sys, json
for l in sys.stdin:
    b = json.loads(l); print(b['n_gpus'], str(b.get('frame_exchange'))[:60], 'value %.2f T' % (b['value']/1e12), 'step %.1f us' % (b['ms_per_step']*1e3), 'e2e %.2f T' % (b['e2e']['value']/1e12), 'e2e step %.1f us' % (b['e2e']['ms_per_step']*1e3), 'count', b['collisions_last_step'])
"
