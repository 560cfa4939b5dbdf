#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py tests/test_gpu_adapter.py -m gpu -x -q 2>&1 | grep -v "^2026" | tail -6
for coop in 0 1; do
OSB200_COOP=$coop timeout 600 python bench.py --no-cpu --steps 3 --warmup 3 > gpurun_out/r2_bench_coop$coop.json 2> gpurun_out/r2_bench_coop$coop.err; echo "coop=$coop rc=$?"
python - $coop <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/r2_bench_coop{sys.argv[1]}.json'))
e=d['extra']
print('step ms',round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()})
print('800x600 ms',round(e['frame_800x600_title']['ms_per_frame'],4),'1080p ms',round(e['frame_1080p']['ms_per_frame'],4),'game', round(e['frame_1080p_game']['ms_per_frame'],4), {k:round(v,4) for k,v in e['frame_1080p_game']['kernel_ms'].items()})
print('rays e2e', e['castray2']['e2e_mrays_s'], e['castray2']['e2e_pageable_mrays_s'])
PY
done
