#!/bin/bash
mkdir -p gpurun_out
for n in 4 8; do
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err ) 2>&1 | grep real
tail -c 300 gpurun_out/bench_${n}gpu.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_${n}gpu.json').read().strip().splitlines()[-1])
print('n_gpus %d value %.1f M/s ms %.4f per_rank %s' % (d['n_gpus'], d['value']/1e6, d['ms_per_step'], [round(x,4) for x in d['per_rank_ms']]))
print('e2e', {k:round(v['value']/1e6,1) for k,v in d['e2e']['modes'].items()}, {k:round(v,1) for k,v in d['e2e']['host_link'].items()})
print({k:round(v['value']/1e6,1) for k,v in d['other_configs'].items()})
PY
done
