#!/bin/bash
mkdir -p gpurun_out
export OSB200_DUAL_STREAM=0 OSB200_STRIP_MB=8192
for so in openspades_b200/lib/var/libosb200_*.so; do
  n=$(basename $so .so); n=${n#libosb200_}
  for v in 32 4096; do
  OSB200_LIB=$PWD/$so python bench.py --no-cpu --no-extra --views $v --steps 4 --warmup 3 > gpurun_out/bench_$n.json 2> gpurun_out/bench_$n.err
  python - "$n" $v <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f'gpurun_out/bench_{n}.json'))
    print(n, sys.argv[2], 'ms',round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()})
except Exception as e:
    print(n,'FAILED',e)
PY
  done
done
