#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/flags_big_check.py 6000 2>&1 | tail -3
timeout 300 python tools/flags_big_check.py 12000 2>&1 | tail -3
