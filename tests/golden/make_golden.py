#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/liboref_*.so, built in place from
/root/reference by oracle/Makefile).  Run in the container that has /root/reference; the fixtures are committed so
that machines without oracle/_ref can still pin the oracle port and the CUDA path.

Only CPU-independent variants are stored: CastRay2 and radiosity are plain IEEE; the terrain frame is the V-IEEE
build (the SSE variants depend on the host CPU's rcpss/rsqrtss and are compared live instead)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracles  # noqa: E402
from openspades_b200 import workloads  # noqa: E402


def main():
    data = open(oracles.TITLE_VXL, "rb").read()
    ref = oracles.Ref("sse2")
    m = ref.map_load_vxl(data)
    solid, color = ref.map_arrays(m)

    # ---- CastRay2: 8192 random + the edge set, three step limits ----
    v0, d = workloads.random_rays(8192, seed=2024)
    e0, ed = workloads.edge_rays()
    keep = np.any(ed != 0, axis=1)
    v0 = np.concatenate([v0, e0[keep]])
    d = np.concatenate([d, ed[keep]])
    hits = {f"hits_{s}": ref.castray2(m, v0, d, s).view(np.uint8).reshape(len(v0), 40) for s in (256, 64, 12)}
    np.savez_compressed(os.path.join(HERE, "castray2_title.npz"), v0=v0, dir=d, **hits)

    # ---- shadow bitmap + radiosity ----
    rr = oracles.RefRadiosity(ref, m, 1)
    bm = rr.shadow_bitmap()
    lo, hi = (248, 250, 24), (263, 257, 39)
    f32 = rr.evaluate_box(lo, hi)
    chunks = {}
    for (cx, cy, cz) in ((16, 16, 2), (0, 31, 0)):
        rr.update_chunk(cx, cy, cz)
        chunks[f"chunk_{cx}_{cy}_{cz}"] = rr.read_chunk(cx, cy, cz)
    rr.close()
    rr2 = oracles.RefRadiosity(ref, m, 2)
    rr2.update_chunk(16, 16, 2)
    chunks["chunk_mode2_16_16_2"] = rr2.read_chunk(16, 16, 2)
    rr2.close()
    np.savez_compressed(os.path.join(HERE, "radiosity_title.npz"), shadow_bitmap=bm, box_lo=np.array(lo),
                        box_hi=np.array(hi), box_f32=f32, **chunks)
    ref.map_free(m)

    # ---- terrain + fog, V-IEEE, title pose 256x256 fovY 68 and one pole-visible pose ----
    refi = oracles.Ref("ieee")
    mi = refi.map_from_arrays(solid, color)
    r = oracles.RefRenderer(refi, 256, 256, -1, 4)
    r.set_map(mi)
    r.set_fog((0.1, 0.1, 0.1), 128.0)
    scenes = np.stack([workloads.title_scene(256, 256, 68.0),
                       workloads.look_at_scene((300.5, 200.5, 8.0), (300.6, 200.9, 9.5), (0, 0, -1), 68.0, 256, 256)])
    fbs, dps = [], []
    for sc in scenes:
        fb, dp = r.render(sc)
        fbs.append(fb & 0xFFFFFF)
        dps.append(dp)
    r.set_map(None)
    r.close()
    refi.map_free(mi)
    np.savez_compressed(os.path.join(HERE, "terrain_ieee_title.npz"), scenes=scenes, bgr=np.stack(fbs),
                        depth=np.stack(dps), fog=np.array([0.1, 0.1, 0.1, 128.0], np.float32))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
