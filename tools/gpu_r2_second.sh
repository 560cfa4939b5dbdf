#!/bin/bash
# round 2, second pass: coherent walker tasks + smaller gather code; parity, bench variants, ncu of the result
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py tests/test_gpu_vxl.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_pytest_terrain2.log; tail -4 gpurun_out/r2_pytest_terrain2.log
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
run base2 A=1
for so in openspades_b200/lib/var/libosb200_*.so; do
  n=$(basename $so .so); n=${n#libosb200_}
  run $n OSB200_LIB=$PWD/$so
done
bash tools/gpu_r2_ncu.sh c
timeout 600 python bench.py --no-cpu > gpurun_out/r2_bench_full2.json 2> gpurun_out/r2_bench_full2.err; echo "full bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_full2.json'))
e=d['extra']; print('1080p ms',round(e['frame_1080p']['ms_per_frame'],4),'game', round(e['frame_1080p_game']['ms_per_frame'],4), {k:round(v,4) for k,v in e['frame_1080p_game']['kernel_ms'].items()})
PY
