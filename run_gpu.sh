#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/gpu_tests.txt
python bench.py > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_r02_reference.json 2> gpurun_out/bench_r02_reference.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read().strip().splitlines()[-1])
print('value %.1f M/s e2e %.1f M/s frac %.3f' % (d['value']/1e6, d['e2e']['value']/1e6, d['roofline']['frac']))
for k,v in d['other_configs'].items(): print(k, '%.1f M/s %.4f ms frac %.3f' % (v['value']/1e6, v['ms_per_step'], v['frac']))
r=json.loads(open('gpurun_out/bench_r02_reference.json').read().strip().splitlines()[-1]); print('reference %.2f M/s' % (r['value']/1e6))
PY
