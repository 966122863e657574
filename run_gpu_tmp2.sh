MG_CHUNKS=1 ncu --set full --clock-control none --import-source on -k regex:mg_obs_kernel -s 215 -c 1 -o gpurun_out/prof_r1k_obs_late -f python bench.py --no-cpu-baseline --steps 200 --warmup 10 > gpurun_out/ncu_o2.log 2>&1
tail -1 gpurun_out/ncu_o2.log | cut -c1-100
