# A/B of GVX_OPT_PDL on one B200: the step-loop parity test first, then bench.py with the option off and on
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_pdl.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pdl.log; tail -n 4 gpurun_out/pytest_pdl.log
for p in 0 1; do
  timeout 120 python bench.py --steps 100 --warmup 3 --no-cpu-baseline --pdl $p > gpurun_out/bench_pdl$p.json 2> gpurun_out/bench_pdl$p.err || tail -n 5 gpurun_out/bench_pdl$p.err
  python -c "
import json; b=json.load(open('gpurun_out/bench_pdl$p.json')); print('pdl $p', b['value'], b['ms_per_step'], b['e2e']['ms_per_step'], b['roofline']['frac'], b['roofline']['kernel_ms_avg'], b['collisions_last_step'])"
done
