# quick N=1 confirmation: all GPU tests, then the bench line
timeout 300 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/pytest_gpu.log
timeout 200 python bench.py --steps 100 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || tail -n 5 gpurun_out/bench_n1.err
python -c "
import json; b=json.load(open('gpurun_out/bench_n1.json')); r=b['roofline']; print(b['value'], b['ms_per_step'], b['e2e']['ms_per_step'], b['e2e']['synchronous_calls']['ms_per_step'], r['frac'], r['kernel_ms_avg'], r['kernel_ms_single_launch_min'], b['collisions_last_step'])"
