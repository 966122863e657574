#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/bench_*.json gpurun_out/bench_*.err
python -m pytest tests -m gpu -x -q -k "arena or crowded or scenarios or fixture" > gpurun_out/gputests.log 2>&1; echo "gpu tests rc=$?" | tee gpurun_out/summary.txt
tail -3 gpurun_out/gputests.log
B="python bench.py --no-cpu-baseline --no-other-configs"
for i in 1 2; do
MG_LIB_PATH=build_variants/lib_before.so $B > gpurun_out/bench_before_$i.json 2> gpurun_out/bench_before_$i.err
$B > gpurun_out/bench_default_$i.json 2> gpurun_out/bench_default_$i.err
done
MG_LIB_PATH=build_variants/lib_timing.so python tools/timing_probe.py 4096 flush 20 stagger > gpurun_out/timing_steady.txt 2>&1
python - <<'PY' | tee -a gpurun_out/summary.txt
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        pk=d['roofline']['per_kernel']; sm=d['step_ms']
        print('%-28s %7.1f M/s  %.4f ms  frac %.3f  step %.4f obs %.4f  free %7.1f  e2e %6.1f/%6.1f  ep %g  min %.4f med %.4f p90 %.4f max %.4f' % (f.split('/')[-1], d['value']/1e6, d['ms_per_step'], d['roofline']['frac'], pk['mg_step_kernel']['ms'], pk['mg_obs_kernel']['ms'], d['free_running']['value']/1e6, d['e2e']['value']/1e6, d['e2e']['modes']['full_rows']['value']/1e6, d['episodes_finished'], sm['min'], sm['median'], sm['p90'], sm['max']))
    except Exception as ex:
        print(f, 'FAILED', ex); print(open(f.replace('.json','.err')).read()[-800:])
PY
