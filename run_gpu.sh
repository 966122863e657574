#!/bin/bash
mkdir -p gpurun_out
MG_LIB_PATH=build_variants/lib_timing.so python tools/handover_probe.py 2>&1 | tail -11 | tee gpurun_out/handover_sliced.txt
MG_LIB_PATH=build_variants/lib_timing.so MG_CHUNKS=1 python tools/handover_probe.py 2>&1 | tail -11 | tee -a gpurun_out/handover_sliced.txt
