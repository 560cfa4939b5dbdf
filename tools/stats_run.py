"""Debug statistics of k_build_lines (library built with -DOSB_STATS into lib/libosb200_stats.so)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["OSB200_LIB"] = os.path.join(ROOT, "openspades_b200", "lib", "libosb200_stats.so")
import openspades_b200 as osb
from openspades_b200 import workloads, binding
from openspades_b200.hostlib import VoxelMap
import torch
m = VoxelMap.load_file(workloads.TITLE_VXL)
nv = int(sys.argv[1]) if len(sys.argv) > 1 else 512
S = workloads.random_scenes(nv, m.solid, 256, 256, seed=4321)
lib = binding.load_library()
with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:
    ctx.map_upload(m.solid.copy(), m.color.copy())
    fb = torch.empty((nv, 256, 256), dtype=torch.int32, device="cuda")
    dp = torch.empty((nv, 256, 256), dtype=torch.float32, device="cuda")
    out = (ctypes.c_ulonglong * 16)()
    lib.osb200_debug_stats(out, 1)
    ctx.render_device(S, 256, 256, (0.8, 1.0, 1.0), 128.0, fb, dp)
    ctx.sync()
    lib.osb200_debug_stats(out, 1)
    names = ["thread-steps", "warp-steps", "tops", "bots", "walls", "warp-span-iters", "nonempty", "land-top", "land-bot", "land-wall"]
    for n, v in zip(names, out):
        print(f"{n:16s} {v:>12d}  per view {v / nv:10.1f}")
