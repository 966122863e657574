#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/gpu_tests.txt
timeout 900 python tools/soak_parity.py 128 10300 2>&1 | tail -2
