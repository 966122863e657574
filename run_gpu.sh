#!/bin/bash
mkdir -p gpurun_out
MG_LIB_PATH=build_variants/lib_timing.so python tools/other_config_split.py 2>&1 | tail -14 | tee gpurun_out/other_split.txt
python tools/other_config_split.py 2>&1 | tail -2
