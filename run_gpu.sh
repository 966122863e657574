#!/bin/bash
mkdir -p gpurun_out
for v in w1 w2 w4; do echo $v; MG_LIB_PATH=build_variants/lib_$v.so python tools/other_config_split.py 2>&1 | tail -2; done
echo default; python tools/other_config_split.py 2>&1 | tail -2
