set -x
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x ) > gpurun_out/pytest_parity.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_parity.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 tests/multi_gpu_check.py > gpurun_out/mgc2.log 2>&1; echo "mgc2 rc=$?"; tail -n 3 gpurun_out/mgc2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 100 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err || tail -n 15 gpurun_out/bench_n2.err
python bench.py --steps 100 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_samebox.json 2>/dev/null
cat gpurun_out/bench_n1_samebox.json gpurun_out/bench_n2.json | python -c "
import sys, json
for l in sys.stdin:
    b = json.loads(l); print(b['n_gpus'], 'value %.2f T' % (b['value']/1e12), 'step %.1f us' % (b['ms_per_step']*1e3), 'e2e %.2f T' % (b['e2e']['value']/1e12), 'e2e step %.1f us' % (b['e2e']['ms_per_step']*1e3), 'sync %.1f us' % (b['e2e']['synchronous_calls']['ms_per_step']*1e3), 'count', b['collisions_last_step'])
"
