#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python tools/soak_parity.py 48 10400 2>&1 | tail -14 | tee gpurun_out/soak.txt
