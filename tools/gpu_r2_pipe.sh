#!/bin/bash
# A/B: k_view_pipeline (OSB200_PIPELINE = window in views) against k_walk_lines + k_gather_fog (0)
mkdir -p gpurun_out
for w in ${WINDOWS:-0 4096 512 256 128 64 0}; do
  OSB200_PIPELINE=$w timeout 180 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_pipe$w.json 2> gpurun_out/r2_bench_pipe$w.err
  echo "rc=$?"
  python - $w <<'PY'
import json,sys
w=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/r2_bench_pipe{w}.json').read().strip().splitlines()[-1])
    print('window',w,'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'checksum', d['result_checksum'], 'parity', d['parity_sample'].get('identical'), d['parity_sample'].get('depth_identical'))
except Exception as e:
    print('window',w,'failed',e); print(open(f'gpurun_out/r2_bench_pipe{w}.err').read()[-1500:])
PY
done
