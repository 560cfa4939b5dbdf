"""Loads Title.vxl through osb200_map_load_vxl and saves it again a few times (the command tools/gpu_final_r01.sh profiles)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openspades_b200 as osb
from openspades_b200 import workloads
data = open(workloads.TITLE_VXL, "rb").read()
with osb.Context(0, osb.NUM_IEEE) as ctx:
    for _ in range(3):
        ctx.map_load_vxl(data, want_arrays=False)
        out = ctx.map_save_vxl()
    ctx.timing_enable(True)
    ctx.timing_read()
    ctx.map_load_vxl(data, want_arrays=False)
    out = ctx.map_save_vxl()
    for k, (n, ms) in ctx.timing_read().items():
        if n:
            print(f"{k:14s} {n:3d} brackets {ms:8.4f} ms")
    print("saved", len(out), "bytes; stream", len(data))
