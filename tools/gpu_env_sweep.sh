#!/bin/bash
# gpu_env_sweep.sh "K=V K2=V2" "K=V" ... : quick bench (views only) once per environment setting
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  i=$((i+1))
  env $cfg python bench.py --no-cpu --no-extra --steps 4 --warmup 3 > gpurun_out/sweep_$i.json 2> gpurun_out/sweep_$i.err
  python - "$cfg" gpurun_out/sweep_$i.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2]))
    print(sys.argv[1] or '(default)', '| Mpx/s',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'sum', d.get('result_checksum'))
except Exception as e:
    print(sys.argv[1],'FAILED',e)
PY
done
