#!/bin/bash
# ncu --set full of the CastRay2 and radiosity kernels (small driver script)
mkdir -p gpurun_out
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
for _ in range(3):
    ctx.castray2_device(dv, dd, 256, out, n)
ctx.sync()
print('chunks', ctx.radiosity_update(-1))
PY
python /tmp/drv.py > gpurun_out/plain_misc.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_castray2' -s 2 -c 2 \
  -o gpurun_out/prof_r01b_misc python /tmp/drv.py > gpurun_out/ncu_misc.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_misc.log
