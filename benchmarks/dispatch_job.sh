# frame dispatch over peer memory: correctness at world N (default 2), then the bench with both fold modes
N=${1:-2}
set -x
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 tests/multi_gpu_check.py > gpurun_out/mgc$N.log 2>&1; rc=$?
echo "mgc$N rc=$rc" >> gpurun_out/mgc$N.log
tail -n 12 gpurun_out/mgc$N.log
[ $rc -ne 0 ] && exit 1
for mode in dispatch multi; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --steps 100 --warmup 3 --fold $mode > gpurun_out/bench_n${N}_$mode.json 2> gpurun_out/bench_n${N}_$mode.err || tail -n 20 gpurun_out/bench_n${N}_$mode.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus $N --steps 100 --warmup 3 --diag > /dev/null 2> gpurun_out/diag_n${N}_dispatch.err
grep "diag rank 0" gpurun_out/diag_n${N}_dispatch.err | cut -c1-400
cat gpurun_out/bench_n${N}_*.json | python -c "
import sys, json
for l in sys.stdin:
    b = json.loads(l); print(b['n_gpus'], b['frame_exchange'][:20], 'value %.2f T' % (b['value']/1e12), 'step %.1f us' % (b['ms_per_step']*1e3), 'e2e %.2f T' % (b['e2e']['value']/1e12), 'e2e step %.1f us' % (b['e2e']['ms_per_step']*1e3), 'count', b['collisions_last_step'])
"
