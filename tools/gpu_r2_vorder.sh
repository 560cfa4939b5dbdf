#!/bin/bash
# A/B: walker takes the views longest-estimated-lines first (OSB200_VIEW_ORDER=1) against index order (0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q 2>&1 | tail -3
for n in ${ORDERS:-0 10 20 40 0 20}; do
  OSB200_VIEW_ORDER=$n timeout 300 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_vorder$n.json 2> gpurun_out/r2_bench_vorder$n.err
  python - $n <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r2_bench_vorder{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print('view order',sys.argv[1],'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'checksum', d['result_checksum'], 'parity', d['parity_sample'].get('identical'), 'launches', d['gpu_launches'])
except Exception as e:
    print('failed', e); print(open(f'gpurun_out/r2_bench_vorder{sys.argv[1]}.err').read()[-1500:])
PY
done
