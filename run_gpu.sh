#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python tools/soak_parity.py 512 10400 2>&1 | tail -14 | tee gpurun_out/soak.txt
