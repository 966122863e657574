for c in 10 12 14 10 12 14 8; do MG_CHUNKS=$c python bench.py --no-cpu-baseline --steps 100 > gpurun_out/bench_c$c.log 2>&1; python - $c <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/bench_c{sys.argv[1]}.log').read().strip().splitlines()[-1]); print('chunks', sys.argv[1], 'M/s %.1f ms/step %.4f free %.1f'%(d['value']/1e6, d['ms_per_step'], d['free_running']['value']/1e6))
PY
done
