#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vecenv.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/gpu_tests.txt
