#!/bin/bash
# full GPU suite + both bench arms + smoke after the castray2 / links / lifetime changes
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -60 | tee gpurun_out/r2_pytest_full.log | tail -8
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/r2_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r2_bench_ref_final.json 2> gpurun_out/r2_bench_ref_final.err; tail -c 600 gpurun_out/r2_bench_ref_final.json
timeout 1200 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; tail -3 gpurun_out/r2_bench_final.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_final.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'traffic',d['roofline']['traffic'],d['roofline'].get('traffic_source'))
print('numerics',d.get('numerics_ms_per_step'))
print('parity',d['parity_sample'])
ex=d['extra']
print('c1',ex['frame_800x600_title']['ms_per_frame'],'c2',ex['frame_1080p']['ms_per_frame'],'game',ex['frame_1080p_game']['ms_per_frame'])
print('rays',{k:v for k,v in ex['castray2'].items() if k!='cpu_baseline' and k!='workload'})
PY
