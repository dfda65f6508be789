# 2-GPU check after the device-only result path: peer-memory checks, then the driver's launch line at N=2
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 tests/multi_gpu_check.py > gpurun_out/mgc2.log 2>&1; echo "mgc2 rc=$?" >> gpurun_out/mgc2.log
tail -n 3 gpurun_out/mgc2.log
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 100 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err || tail -n 5 gpurun_out/bench_n2.err
python -c "
import json; b=json.load(open('gpurun_out/bench_n2.json')); print(b['n_gpus'], b['value'], b['ms_per_step'], b['e2e']['ms_per_step'], b['collisions_last_step'])"
