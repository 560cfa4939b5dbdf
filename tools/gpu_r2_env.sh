#!/bin/bash
# A/B of an environment knob: ENVV=NAME, VALUES="a b a b"; headline bench without extras
mkdir -p gpurun_out
for v in $VALUES; do
  env $ENVV=$v timeout 300 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_env_$v.json 2> gpurun_out/r2_bench_env_$v.err
  python - $v <<'PY'
import json,sys,os
try:
    d=json.loads(open(f'gpurun_out/r2_bench_env_{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print(os.environ.get('ENVV'),sys.argv[1],'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'checksum', d['result_checksum'], 'parity', d['parity_sample'].get('identical'))
except Exception as e:
    print(sys.argv[1],'failed', e); print(open(f'gpurun_out/r2_bench_env_{sys.argv[1]}.err').read()[-1200:])
PY
done
