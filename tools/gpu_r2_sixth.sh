#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py -m gpu -x -q 2>&1 | grep -v "^2026" | tail -4
timeout 600 python bench.py --no-cpu --steps 3 --warmup 3 > gpurun_out/r2_bench_path.json 2> gpurun_out/r2_bench_path.err; echo "rc=$?"
python - <<'PY'
import json,sys
d=json.load(open('gpurun_out/r2_bench_path.json'))
e=d['extra']
print('step ms',round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()})
print('800x600 ms',round(e['frame_800x600_title']['ms_per_frame'],4),'1080p ms',round(e['frame_1080p']['ms_per_frame'],4),'game', round(e['frame_1080p_game']['ms_per_frame'],4), {k:round(v,4) for k,v in e['frame_1080p_game']['kernel_ms'].items()})
PY
