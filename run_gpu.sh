#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
python bench.py --no-cpu-baseline --no-other-configs --steps 100 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('arena %.1f M/s %.4f ms free %.1f' % (d['value']/1e6, d['ms_per_step'], d['free_running']['value']/1e6), d['step_ms'])"
done
