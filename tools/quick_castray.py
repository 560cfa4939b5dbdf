"""Throw-away timing of the CastRay2 kernel (device-resident inputs), CUDA events on the ctx stream."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

n = 1 << 24
m = VoxelMap.load_file(workloads.TITLE_VXL)
ctx = osb.Context(0, osb.NUM_IEEE)
ctx.map_upload(m.solid, m.color)
v0, d = workloads.random_rays(n, 1234)
dv = torch.from_numpy(v0).cuda(); dd = torch.from_numpy(d).cuda()
out = torch.empty(n * 40, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
st = torch.cuda.ExternalStream(ctx.stream)
for steps in (256, 64):
    with torch.cuda.stream(st):
        for _ in range(3):
            ctx.castray2_device(dv, dd, steps, out, n)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(5):
            ctx.castray2_device(dv, dd, steps, out, n)
        e1.record(st)
    ctx.sync(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"castray2 n={n} maxSteps={steps}: {ms:.3f} ms  {n/ms/1e3:.1f} Mrays/s  {n*64/ms/1e6:.1f} GB/s algorithmic")
