#!/bin/bash
# runs the quick bench once per library variant under openspades_b200/lib/var/
mkdir -p gpurun_out
for so in openspades_b200/lib/var/libosb200_*.so; do
  n=$(basename $so .so); n=${n#libosb200_}
  OSB200_LIB=$PWD/$so python bench.py --no-cpu --no-extra --steps 4 --warmup 3 > gpurun_out/bench_$n.json 2> gpurun_out/bench_$n.err
  python - "$n" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f'gpurun_out/bench_{n}.json'))
    print(n, 'Mpx/s',round(d['value']),'ms',round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'sum', d.get('result_checksum'))
except Exception as e:
    print(n,'FAILED',e)
PY
done
