#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/bench_*.json gpurun_out/bench_*.err
B="timeout 300 python bench.py --no-cpu-baseline --no-other-configs --steps 100"
run() { name=$1; shift; env "$@" $B > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; }
run a1 A=1; run a2 A=1; run a3 A=1
python - <<'PY' | tee gpurun_out/summary.txt
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        pk=d['roofline']['per_kernel']; sm=d['step_ms']
        print('%-28s %7.1f M/s  %.4f ms  frac %.3f  step %.4f obs %.4f free %7.1f  min %.4f med %.4f p90 %.4f max %.4f' % (f.split('/')[-1], d['value']/1e6, d['ms_per_step'], d['roofline']['frac'], pk['mg_step_kernel']['ms'], pk['mg_obs_kernel']['ms'], d['free_running']['value']/1e6, sm['min'], sm['median'], sm['p90'], sm['max']))
    except Exception as ex:
        print(f, 'FAILED', ex); print(open(f.replace('.json','.err')).read()[-500:])
PY
