#!/bin/bash
# terrain / renderer / models parity tests, then two headline bench runs without extras
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py tests/test_gpu_adapter.py -m gpu -x -q 2>&1 | tail -4
for i in 1 2; do
  timeout 300 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_quick$i.json 2> gpurun_out/r2_bench_quick$i.err
  python - $i <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r2_bench_quick{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print('step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'checksum', d['result_checksum'], 'parity', d['parity_sample'].get('identical'), d['parity_sample'].get('depth_identical'))
except Exception as e:
    print('failed', e); print(open(f'gpurun_out/r2_bench_quick{sys.argv[1]}.err').read()[-1500:])
PY
done
