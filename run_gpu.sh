#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/bench_*.json gpurun_out/bench_*.err
python tools/obs_ab.py 2>&1 | tail -1 | tee gpurun_out/obs_ab.txt
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:mg_obs_kernel -c 1 -o gpurun_out/r02_obs -f python tools/profile_steady.py > gpurun_out/ncu_obs.log 2>&1
B="python bench.py --no-cpu-baseline --no-other-configs --steps 100"
$B > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
MG_STEP_MODE=queue MG_QUEUE_EARLY=8 MG_QUEUE_LATE=0 $B > gpurun_out/bench_queue_e8_l0.json 2> gpurun_out/bench_queue_e8_l0.err
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
