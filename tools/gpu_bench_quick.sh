#!/bin/bash
# quick A/B: bench only (no cpu leg), optional env passthrough
mkdir -p gpurun_out
for v in "$@"; do export "$v"; done
python bench.py --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']))
print({k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()}, round(d['roofline']['frac'],4))
e=d['extra']; print('1080p ms',round(e['frame_1080p']['ms_per_frame'],4),'rays ms',round(e['castray2']['ms_per_batch'],3), 'Mrays/s', round(e['castray2']['mrays_s']))
PY
tail -3 gpurun_out/bench.err
