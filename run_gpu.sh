#!/bin/bash
mkdir -p gpurun_out
MG_LIB_PATH=build_variants/lib_timing.so MG_CHUNKS=1 python tools/timing_probe.py 4096 flush 12 stagger > gpurun_out/timing_default.txt 2>&1
tail -10 gpurun_out/timing_default.txt
