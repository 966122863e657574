#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/gpu_tests.txt
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:mg_step_kernel -c 1 -o gpurun_out/r02c_step -f python tools/profile_steady.py > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log
