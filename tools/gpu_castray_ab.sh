#!/bin/bash
cat > /tmp/cr.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap
m = VoxelMap.load_file(workloads.TITLE_VXL)
ctx = osb.Context(0, osb.NUM_IEEE)
ctx.map_upload(m.solid.copy(), m.color.copy())
n = 1 << 24
v0, d = workloads.random_rays(n, seed=4321)
d /= np.linalg.norm(d, axis=1, keepdims=True)
dv, dd = torch.from_numpy(v0).cuda(), torch.from_numpy(d.astype(np.float32)).cuda()
out = torch.empty((n, 4), dtype=torch.int32, device='cuda')
res = []
for L in (128.0, 16.0):
    ctx.castray_device(dv, dd, L, out, n); ctx.sync()
    ctx.timing_enable(True); ctx.timing_read()
    for _ in range(5):
        ctx.castray_device(dv, dd, L, out, n)
    ctx.sync()
    res.append(ctx.timing_read()["k_castray"][1] / 5)
    ctx.timing_enable(False)
print("len128 %.3f ms   len16 %.3f ms   checksum %d" % (res[0], res[1], int(out.sum().item())))
PY
echo "default"; python /tmp/cr.py
for v in openspades_b200/lib/var/libosb200_*.so; do echo "$v"; OSB200_LIB=$PWD/$v python /tmp/cr.py; done
