#!/bin/bash
python -m pytest tests/test_gpu_ambient.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
python - <<'PY'
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
for L in (128.0, 16.0, 512.0):
    ctx.castray_device(dv, dd, L, out, n); ctx.sync()
    ctx.timing_enable(True); ctx.timing_read()
    for _ in range(5):
        ctx.castray_device(dv, dd, L, out, n)
    ctx.sync()
    ms = ctx.timing_read()["k_castray"][1] / 5
    print(f"length {L}: {ms:.3f} ms = {n / ms / 1e6:.2f} Grays/s, hit {float((out[:,0]!=0).float().mean()):.3f}")
    ctx.timing_enable(False)
PY
