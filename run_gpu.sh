#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/bench_*.json gpurun_out/bench_*.err
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "queue" 2>&1 | tail -3 | tee gpurun_out/gpu_tests.txt
B="python bench.py --no-cpu-baseline --no-other-configs --steps 100"
$B > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
run() { name=$1; shift; env "$@" MG_STEP_MODE=queue $B > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; }
run q24_e8 MG_QUEUE_EARLY=8 MG_QUEUE_LATE=0
run q24_e4l4 MG_QUEUE_EARLY=4 MG_QUEUE_LATE=4
for v in q28 q28r4 q28r1; do
  run ${v}_e8 MG_LIB_PATH=build_variants/lib_$v.so MG_QUEUE_EARLY=8 MG_QUEUE_LATE=0
done
run q28_e6 MG_LIB_PATH=build_variants/lib_q28.so MG_QUEUE_EARLY=6 MG_QUEUE_LATE=0
run q28_e4l4 MG_LIB_PATH=build_variants/lib_q28.so MG_QUEUE_EARLY=4 MG_QUEUE_LATE=4
run q28_e0l8 MG_LIB_PATH=build_variants/lib_q28.so MG_QUEUE_EARLY=0 MG_QUEUE_LATE=8
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
