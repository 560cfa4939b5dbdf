#!/bin/bash
# ncu --set full of the two terrain kernels on one chunk-sized batch (680 views = one strip-budget chunk)
mkdir -p gpurun_out
CMD="python bench.py --views 680 --steps 2 --warmup 1 --no-extra --no-cpu"
$CMD > gpurun_out/plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_build_lines|k_gather_fog' -s 2 -c 2 \
  -o gpurun_out/prof_r01c_terrain $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
