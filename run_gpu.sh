#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/gpu_tests.txt
python bench.py > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_r02_reference.json 2> gpurun_out/bench_r02_reference.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read().strip().splitlines()[-1])
print('value %.1f M/s ms %.4f frac %.3f e2e %.1f' % (d['value']/1e6, d['ms_per_step'], d['roofline']['frac'], d['e2e']['value']/1e6), {k:round(v['value']/1e6,1) for k,v in d['e2e']['modes'].items()})
print({k:round(v['value']/1e6,1) for k,v in d['other_configs'].items()}, 'cpu', d['cpu_baseline']['value'])
r=json.loads(open('gpurun_out/bench_r02_reference.json').read().strip().splitlines()[-1]); print('reference %.2f M/s' % (r['value']/1e6))
PY
