#!/bin/bash
# CastRay2 round: tests, then A/B of the DDA steps between two idle checks (2 / 4 / 8), then e2e with 2..8 copy threads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_castray2.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2_rays_pytest.log
for v in base spc8 spc2 base; do
  if [ $v = base ]; then unset OSB200_LIB; else export OSB200_LIB=$PWD/openspades_b200/lib/var/libosb200_$v.so; fi
  echo "== $v"; timeout 300 python tools/ray_bench.py 2>&1 | tail -1 | tee gpurun_out/r2_rays_$v.json
done
unset OSB200_LIB
for t in 2 4 8 12; do echo "== copy threads $t"; OSB200_COPY_THREADS=$t timeout 300 python tools/ray_bench.py 2>&1 | tail -1 | tee gpurun_out/r2_rays_threads$t.json; done
timeout 120 python tools/d2h_ceiling.py 2>&1 | tail -1
nproc; lscpu | grep -E "Model name|NUMA|Socket" 
