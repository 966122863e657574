#!/bin/bash
# One GPU session: tests, benches of the launch modes and build variants, timing probe. Output under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/gputests.log 2>&1; echo "gpu tests rc=$?" | tee -a gpurun_out/summary.txt
tail -5 gpurun_out/gputests.log
python bench.py > gpurun_out/bench_sliced_full.json 2> gpurun_out/bench_sliced_full.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
B="python bench.py --no-cpu-baseline --no-other-configs"
MG_STEP_MODE=fused $B > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err
MG_STEP_MODE=queue $B > gpurun_out/bench_queue.json 2> gpurun_out/bench_queue.err
for v in rows2 rows8 w8c3 statred; do
  MG_STEP_MODE=fused MG_LIB_PATH=build_variants/lib_$v.so $B > gpurun_out/bench_fused_$v.json 2> gpurun_out/bench_fused_$v.err
done
MG_LIB_PATH=build_variants/lib_statred.so $B > gpurun_out/bench_sliced_statred.json 2> gpurun_out/bench_sliced_statred.err
MG_STEP_MODE=fused python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs > gpurun_out/bench_fused_20.json 2> gpurun_out/bench_fused_20.err
MG_CHUNKS=1 MG_LIB_PATH=build_variants/lib_timing.so python tools/timing_probe.py 4096 flush 20 stagger > gpurun_out/timing_steady.txt 2>&1
python - <<'PY' | tee -a gpurun_out/summary.txt
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        pk=d['roofline']['per_kernel']
        print('%-34s %7.1f M/s  %.4f ms  frac %.3f  step %.4f obs %.4f  free %7.1f  e2e %6.1f / full %6.1f  ep %g  p90 %.4f max %.4f' % (f.split('/')[-1], d['value']/1e6, d['ms_per_step'], d['roofline']['frac'], pk['mg_step_kernel']['ms'], pk['mg_obs_kernel']['ms'], d['free_running']['value']/1e6, d['e2e']['value']/1e6, d['e2e']['modes']['full_rows']['value']/1e6, d['episodes_finished'], d['step_ms']['p90'], d['step_ms']['max']))
    except Exception as ex:
        print(f, 'FAILED', ex); print(open(f.replace('.json','.err')).read()[-800:])
PY
