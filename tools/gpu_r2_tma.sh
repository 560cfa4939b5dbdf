#!/bin/bash
# A/B: line records of the gather staged in shared memory by one bulk asynchronous copy per CTA (cp.async.bulk + mbarrier)
mkdir -p gpurun_out
OSB200_LIB=$PWD/openspades_b200/lib/var/libosb200_tma.so timeout 900 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q 2>&1 | tail -3
for n in base tma base tma; do
  if [ $n = tma ]; then export OSB200_LIB=$PWD/openspades_b200/lib/var/libosb200_tma.so; else unset OSB200_LIB; fi
  timeout 600 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_ab_$n.json 2> gpurun_out/r2_bench_ab_$n.err
  python - $n <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/r2_bench_ab_{sys.argv[1]}.json'))
print(sys.argv[1],'step ms',round(d['ms_per_step'],3), {k:round(v,4) for k,v in d['roofline']['kernel_ms_per_step'].items()})
PY
done
