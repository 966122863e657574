export MG_STEP_MODE=queue MG_QUEUE_EARLY=8 MG_QUEUE_LATE=0
MG_LIB_PATH=build_variants/lib_q32.so timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_q32.log 2>&1
tail -c 300 gpurun_out/bench_q32.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_q32.log').read().strip().splitlines()[-1])
print('queue q32 M/s %.1f ms/step %.4f'%(d['value']/1e6, d['ms_per_step']))
PY
