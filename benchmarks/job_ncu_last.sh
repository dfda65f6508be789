# one --set full capture of the default collide after the tail changes (bench.py ran plain in the calls before)
timeout 24 ncu --set full --clock-control none --import-source on -k regex:k_collide_summary2 -s 6 -c 1 -f -o gpurun_out/r1_collide_summary2_pdl python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_collide_s2.log 2>&1; echo "rc=$?"
