#!/bin/bash
# usage: tools/bench_variants.sh name ... ; runs bench.py against build_variants/lib_<name>.so and prints the key numbers
for v in "$@"; do
  if [ "$v" = default ]; then python bench.py --no-cpu-baseline > gpurun_out/bench_$v.log 2>&1; else MG_LIB_PATH=build_variants/lib_$v.so python bench.py --no-cpu-baseline > gpurun_out/bench_$v.log 2>&1; fi
  python - "$v" <<'PY'
import json,sys
v=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/bench_{v}.log').read().strip().splitlines()[-1])
    print(v, 'M/s %.1f ms/step %.4f obs_ms %.4f step_ms %.4f free %.1f e2e %.1f (d2h %.2f MB) full %.1f' % (d['value']/1e6, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['step_kernel_ms'], d['free_running']['value']/1e6, d['e2e']['value']/1e6, d['e2e']['d2h_bytes_per_step']/1e6, d['e2e']['full_rows']['value']/1e6))
except Exception as ex:
    print(v, 'FAILED', ex); print(open(f'gpurun_out/bench_{v}.log').read()[-600:])
PY
done
