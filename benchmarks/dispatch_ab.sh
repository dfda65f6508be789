N=${1:-2}
for v in coop plain coop plain; do
  [ $v = plain ] && export GVX_DISPATCH_PLAIN_LAUNCH=1 || unset GVX_DISPATCH_PLAIN_LAUNCH
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus $N --steps 200 --warmup 3 --diag > /dev/null 2> gpurun_out/diag_ab.err
  echo "$v: $(grep -o 'diag rank 0: {[^}]*}' gpurun_out/diag_ab.err)"
done
