#!/bin/bash
# Round-1 profile set: launch list + ncu --set full of the terrain kernels at the bench's launch size, then the
# ray / radiosity / ambient kernels.  Every ncu pass runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --views 4096 --steps 2 --warmup 1 --no-extra --no-cpu"
$CMD > gpurun_out/plain_r01.json 2> gpurun_out/plain_r01.err || { echo "plain run failed"; tail -5 gpurun_out/plain_r01.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_launches_views4096.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_build_lines|k_gather_fog' -s 2 -c 2 \
  -o gpurun_out/prof_r01d_terrain $CMD > gpurun_out/ncu_full.log 2>&1
echo "terrain full rc=$?"
cat > /tmp/drv.py <<'PY'
import sys, numpy as np
sys.path.insert(0, '/root/repo')
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap
import torch
m = VoxelMap.load_file(workloads.TITLE_VXL)
ctx = osb.Context(0, osb.NUM_IEEE)
ctx.map_upload(m.solid.copy(), m.color.copy())
n = 1 << 22
v0, d = workloads.random_rays(n, seed=1234)
dv, dd = torch.from_numpy(v0).cuda(), torch.from_numpy(d).cuda()
out = torch.empty((n, 10), dtype=torch.int32, device='cuda')
out4 = torch.empty((n, 4), dtype=torch.int32, device='cuda')
du = torch.from_numpy((d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)).cuda()
for _ in range(3):
    ctx.castray2_device(dv, dd, 256, out, n)
    ctx.castray_device(dv, du, 128.0, out4, n)
ctx.sync()
print('rad chunks', ctx.radiosity_update(-1))
print('ao chunks', ctx.ambient_update(-1))
ctx.sync()
PY
python /tmp/drv.py > gpurun_out/plain_misc.log 2>&1 || { echo "misc plain failed"; tail -5 gpurun_out/plain_misc.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'k_castray2|k_castray|k_radiosity|k_ao_chunk' -s 2 -c 6 \
  -o gpurun_out/prof_r01d_misc python /tmp/drv.py > gpurun_out/ncu_misc.log 2>&1
echo "misc full rc=$?"; tail -2 gpurun_out/ncu_misc.log
