MG_CHUNKS=1 python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/plain2.log 2>&1 && MG_CHUNKS=1 ncu --set full --clock-control none --import-source on -k regex:mg_step_kernel -s 6 -c 1 -o gpurun_out/prof_r1h_step -f python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/ncu_s.log 2>&1
tail -1 gpurun_out/ncu_s.log | cut -c1-100
