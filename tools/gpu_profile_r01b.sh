#!/bin/bash
# Round-1 closing profile set (after the strip-mask change): full GPU suite, both bench arms, launch list + ncu --set full of the
# terrain kernels at the bench's launch size, k_castray (refill) and k_links_flood.  ncu only after the plain run exited 0.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference > gpurun_out/bench_ref_r01j.json 2> gpurun_out/bench_ref_r01j.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench_r01j.json 2> gpurun_out/bench_r01j.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r01j.err
CMD="python bench.py --views 4096 --steps 2 --warmup 1 --no-extra --no-cpu"
$CMD > gpurun_out/plain_r01j.json 2> gpurun_out/plain_r01j.err || { echo "plain run failed"; tail -5 gpurun_out/plain_r01j.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_launches_views4096.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_build_lines|k_gather_fog' -s 2 -c 2 \
  -o gpurun_out/prof_r01j_terrain $CMD > gpurun_out/ncu_full.log 2>&1
echo "terrain full rc=$?"
cat > /tmp/drv2.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap
import vxl_cases
m = VoxelMap.load_file(workloads.TITLE_VXL)
ctx = osb.Context(0, osb.NUM_IEEE)
ctx.map_upload(m.solid.copy(), m.color.copy())
n = 1 << 22
v0, d = workloads.random_rays(n, seed=1234)
du = torch.from_numpy((d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)).cuda()
dv = torch.from_numpy(v0).cuda()
out4 = torch.empty((n, 4), dtype=torch.int32, device='cuda')
for _ in range(3):
    ctx.castray_device(dv, du, 128.0, out4, n)
ctx.sync()
s, c = vxl_cases.synthetic_arrays(2)
ctx.map_upload(s, c)
print('cave map flood launches', ctx.links_rebuild())
ctx.minimap()
PY
python /tmp/drv2.py > gpurun_out/plain_misc2.log 2>&1 || { echo "misc plain failed"; tail -5 gpurun_out/plain_misc2.log; exit 1; }
cat gpurun_out/plain_misc2.log
ncu --set full --clock-control none --import-source on -k regex:'k_castray|k_links_flood|k_minimap' -s 2 -c 4 \
  -o gpurun_out/prof_r01j_misc python /tmp/drv2.py > gpurun_out/ncu_misc2.log 2>&1
echo "misc full rc=$?"; tail -2 gpurun_out/ncu_misc2.log
