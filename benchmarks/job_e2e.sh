set -x
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x ) > gpurun_out/pytest_parity.log 2>&1; echo "pytest rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -5 gpurun_out/pytest_parity.log; tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n1.json').read().strip().splitlines()[-1])
print('value', d['value'], d['ms_per_step']); print('e2e', {k:v for k,v in d['e2e'].items() if k not in ('note','mode')})
PY
