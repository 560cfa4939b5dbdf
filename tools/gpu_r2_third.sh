#!/bin/bash
# round 2, third pass: staging sync fix, gather batch / occupancy variants, chunk overlap experiments
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_terrain.py -m gpu -x -q 2>&1 | tail -4
run() { # name, env...
  n=$1; shift
  env "$@" timeout 600 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_$n.json 2> gpurun_out/r2_bench_$n.err
  python - "$n" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f'gpurun_out/r2_bench_{n}.json'))
    print(n, 'ms',round(d['ms_per_step'],3), 'e2e ms', round(d['e2e']['ms_per_step'],2), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()}, 'sm', d['clocks']['sm_mhz'])
except Exception as e:
    print(n,'FAILED',e)
PY
}
run base3 A=1
for so in openspades_b200/lib/var/libosb200_*.so; do
  n=$(basename $so .so); n=${n#libosb200_}
  run $n OSB200_LIB=$PWD/$so
done
for ch in 2 4 8; do for wb in 0 6 5 4; do
  run ch${ch}_wb${wb} OSB200_MIN_CHUNKS=$ch OSB200_WALK_BLOCKS=$wb
done; done
