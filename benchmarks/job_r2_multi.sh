#!/bin/bash
# round 2, multi-GPU job: N = $1 GPUs of one box. multi_gpu_check (kept log), SPMD drop-in check, bench at N (and the
# smaller N's that fit the same box when $2 = "scale")
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
[ "$3" = "nocheck" ] || timeout 600 $TR --nproc-per-node $N --master-port 29511 tests/multi_gpu_check.py > gpurun_out/m_check_n$N.log 2>&1
echo "multi_gpu_check rc=$?" | tee -a gpurun_out/m_check_n$N.log
tail -3 gpurun_out/m_check_n$N.log
[ "$3" = "nocheck" ] || timeout 300 python benchmarks/shard_spmd_check.py $N > gpurun_out/m_spmd_n$N.log 2>&1
echo "shard_spmd_check rc=$?" | tee -a gpurun_out/m_spmd_n$N.log
tail -3 gpurun_out/m_spmd_n$N.log
LIST="$N"; [ "$2" = "scale" ] && LIST="1 2 4 8"; [ -n "$2" ] && [ "$2" != "scale" ] && LIST="$2"
for K in $LIST; do
  [ $K -gt $N ] && continue
  if [ $K -eq 1 ]; then CMD="python bench.py --gpus 1"; else CMD="$TR --nproc-per-node $K --master-port 2952$K bench.py --gpus $K"; fi
  timeout 900 $CMD --steps 20 --warmup 3 > gpurun_out/m_bench_n$K.json 2> gpurun_out/m_bench_n$K.err
  echo "bench N=$K rc=$?"; tail -c 600 gpurun_out/m_bench_n$K.err
  python - <<PY
import json
try:
    l = json.loads(open("gpurun_out/m_bench_n$K.json").read().strip().splitlines()[-1])
    print("N=$K value %.3e ms %.3f e2e %.3e default %.3e fold %s c5 %s" % (l["value"], l["ms_per_step"], l["e2e"]["value"],
          l["default_path"]["value"], l["insert"]["how"], {k: (v["ms_per_step"], v["parity"]) for k, v in l["c5"].items() if isinstance(v, dict)}))
except Exception as e:
    print("N=$K no line:", e)
PY
done
