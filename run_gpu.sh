#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vecenv.py -m gpu -x -q 2>&1 | tail -2
python tools/profile_steady.py > gpurun_out/profile_plain.log 2>&1 || exit 1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches.csv python tools/profile_steady.py > gpurun_out/ncu_launches.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:mg_obs_kernel -c 1 -o gpurun_out/r02_obs -f python tools/profile_steady.py > gpurun_out/ncu_obs.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:mg_step_kernel -c 1 -o gpurun_out/r02_step -f python tools/profile_steady.py > gpurun_out/ncu_step.log 2>&1
python bench.py > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_r02_reference.json 2> gpurun_out/bench_r02_reference.err
tail -c 300 gpurun_out/bench_r02.err
