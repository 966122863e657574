python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.log').read().strip().splitlines()[-1])
print(d['value']/1e6, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['step_kernel_ms'], d['free_running']['value']/1e6, d['e2e']['value']/1e6)
PY
