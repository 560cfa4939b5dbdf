#!/usr/bin/env python
"""CastRay2 A/B helper: device-timed batch and end-to-end (pinned / pageable caller buffers) for the library that
OSB200_LIB points at (default: the in-tree build).  Prints one JSON line."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import openspades_b200 as osb  # noqa: E402
from openspades_b200 import workloads  # noqa: E402
from openspades_b200.hostlib import VoxelMap  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
m = VoxelMap.load_file(workloads.TITLE_VXL)
v0, d = workloads.random_rays(n, seed=1234)
dev = torch.device("cuda", 0)
out = {"rays": n}
with osb.Context(device=0, numerics=osb.NUM_IEEE) as ctx:
    ctx.map_upload(m.solid.copy(), m.color.copy())
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    dv, dd = torch.from_numpy(v0).to(dev), torch.from_numpy(d).to(dev)
    do = torch.zeros(n * 10, dtype=torch.int32, device=dev)
    for _ in range(3):
        ctx.castray2_device(dv, dd, 256, do, n)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(5):
        ctx.castray2_device(dv, dd, 256, do, n)
    e1.record(ext)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / 5
    out["device_ms"] = ms
    out["device_mrays_s"] = n / ms / 1e3
    out["checksum"] = int(do.to(torch.int64).sum().item())
    # end to end, pageable numpy buffers
    res = np.zeros(n, osb.RAYHIT_DTYPE)
    ctx.castray2(v0, d, 256, out=res)
    t0 = time.perf_counter()
    for _ in range(3):
        ctx.castray2(v0, d, 256, out=res)
    out["e2e_pageable_mrays_s"] = 3 * n / (time.perf_counter() - t0) / 1e6
    pv, pd = torch.from_numpy(v0).pin_memory(), torch.from_numpy(d).pin_memory()
    po = torch.zeros(n * 40, dtype=torch.uint8).pin_memory()
    ctx.castray2(pv, pd, 256, out=po)
    t0 = time.perf_counter()
    for _ in range(3):
        ctx.castray2(pv, pd, 256, out=po)
    out["e2e_pinned_mrays_s"] = 3 * n / (time.perf_counter() - t0) / 1e6
    out["pinned_equals_pageable"] = bool(po.numpy().tobytes() == res.tobytes())
print(json.dumps(out))
