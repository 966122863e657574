#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/gpu_tests.txt
python tools/other_config_split.py 2>&1 | tail -2
python bench.py --no-cpu-baseline --no-other-configs --steps 100 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('arena %.1f M/s' % (d['value']/1e6))"
