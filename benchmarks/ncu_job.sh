set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python benchmarks/profile_kernels.py > gpurun_out/ncu_plain_kernels.log 2>&1 || exit 2
ncu --set full --clock-control none --import-source on -k regex:k_collide_list_map -s 1 -c 2 -f -o gpurun_out/r1_list python benchmarks/profile_kernels.py list > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_batch_collide|k_env_bitmap' -s 2 -c 2 -f -o gpurun_out/r1_batch python benchmarks/profile_kernels.py batch > gpurun_out/ncu_batch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_insert|k_update_logodds' -s 2 -c 2 -f -o gpurun_out/r1_insert python benchmarks/profile_kernels.py insert > gpurun_out/ncu_insert.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_collide_voxelwise -s 4 -c 1 -f -o gpurun_out/r1_collide python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_collide.log 2>&1
ls -la gpurun_out/
