#!/bin/bash
# A/B: walker tasks queued longest-estimated-first (OSB200_TASK_ORDER=1, default) against view order (0)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q 2>&1 | tail -3
for n in 0 1 0 1; do
  OSB200_TASK_ORDER=$n timeout 600 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_order$n.json 2> gpurun_out/r2_bench_order$n.err
  python - $n <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/r2_bench_order{sys.argv[1]}.json'))
print('order',sys.argv[1],'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'parity', d['parity_sample'])
PY
done
