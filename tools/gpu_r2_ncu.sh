#!/bin/bash
# round 2: bench line of the current build, then ncu --set full (+ source) of the terrain kernels on a 1024-view launch
mkdir -p gpurun_out
tag=${1:-a}
timeout 600 python bench.py --no-cpu --no-extra --steps 5 --warmup 3 > gpurun_out/r2_bench_$tag.json 2> gpurun_out/r2_bench_$tag.err
python - "$tag" <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/r2_bench_{sys.argv[1]}.json'))
print('ms',round(d['ms_per_step'],3), 'e2e ms', round(d['e2e']['ms_per_step'],2), {k:round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()})
PY
CMD="python bench.py --views 1024 --steps 2 --warmup 1 --no-extra --no-cpu"
$CMD > gpurun_out/r2_plain_$tag.json 2> gpurun_out/r2_plain_$tag.err || { echo "plain run failed"; tail -5 gpurun_out/r2_plain_$tag.err; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_walk_lines|k_gather_fog|k_line_setup' -s 3 -c 3 \
  -o gpurun_out/r2_prof_$tag $CMD > gpurun_out/r2_ncu_$tag.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r2_ncu_$tag.log; ls -la gpurun_out/r2_prof_$tag.ncu-rep
