#!/bin/bash
# k_view_pipeline: roles per SM (OSB200_PIPE_SPLIT = walker SMs per 16)
mkdir -p gpurun_out
for cfg in "0 0" "4096 0" "4096 10" "4096 9" "4096 11" "512 10" "256 10"; do
  set -- $cfg
  OSB200_PIPELINE=$1 OSB200_PIPE_SPLIT=$2 timeout 180 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_pipe$1_s$2.json 2> gpurun_out/r2_bench_pipe$1_s$2.err
  python - $1 $2 <<'PY'
import json,sys
w,sp=sys.argv[1:3]
try:
    d=json.loads(open(f'gpurun_out/r2_bench_pipe{w}_s{sp}.json').read().strip().splitlines()[-1])
    print('window',w,'split',sp,'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'checksum', d['result_checksum'], 'parity', d['parity_sample'].get('identical'))
except Exception as e:
    print('window',w,'failed',e); print(open(f'gpurun_out/r2_bench_pipe{w}_s{sp}.err').read()[-800:])
PY
done
