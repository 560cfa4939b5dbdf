#!/bin/bash
# kernel time vs views per launch (single chunk, single stream): exposes the fixed per-launch cost
mkdir -p gpurun_out
export OSB200_DUAL_STREAM=0 OSB200_STRIP_MB=8192
for n in 32 64 128 256 512 1024 2048 4096; do
  python bench.py --no-cpu --no-extra --views $n --steps 6 --warmup 3 > gpurun_out/views_$n.json 2> gpurun_out/views_$n.err
  python - $n <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f'gpurun_out/views_{n}.json'))
    print(n, 'ms',round(d['ms_per_step'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, d['roofline']['kernel_launches_per_step'])
except Exception as e:
    print(n,'FAILED',e)
PY
done
