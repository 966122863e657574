#!/bin/bash
mkdir -p gpurun_out
for v in o_new o_st16 o_nofg o_late o_old o_new o_old; do MG_LIB_PATH=build_variants/lib_$v.so python tools/obs_ab.py 2>&1 | tail -1; done | tee gpurun_out/obs_ab.txt
