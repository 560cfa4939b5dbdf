#!/usr/bin/env python
"""Generates tests/golden/ambient_title.npz from the UNMODIFIED reference (oracle/_ref/liboref_sse2.so): GameMap::CastRay
records and GLAmbientShadowRenderer Evaluate / UpdateChunk outputs on Title.vxl, together with the 16 random ray
directions of the instance that produced them.  Run where /root/reference exists; the fixture is committed."""
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
    ref = oracles.Ref("sse2")
    m = ref.map_load_vxl(open(oracles.TITLE_VXL, "rb").read())
    v0, d = workloads.random_rays(8192, seed=2025)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    v0[:1024, 2] -= 40.0
    v0[1024:2048, 2] += 30.0
    e0, ed = workloads.edge_rays()
    v0 = np.concatenate([v0, e0]).astype(np.float32)
    d = np.concatenate([d.astype(np.float32), ed])
    hits = {f"hits_{int(L * 10)}": ref.castray(m, v0, d, L) for L in (16.0, 128.0, 3.5)}
    ao = oracles.RefAmbient(ref, m)
    lo, hi = (248, 250, 24), (263, 257, 39)
    box = ao.evaluate_box(lo, hi)
    chunks = {}
    for c in ((16, 16, 2), (0, 31, 0), (31, 0, 3)):
        ao.update_chunk(*c)
        chunks["chunk_%d_%d_%d" % c] = ao.read_chunk(*c)
    np.savez_compressed(os.path.join(HERE, "ambient_title.npz"), v0=v0, dir=d, rays=ao.rays(), box_lo=np.array(lo),
                        box_hi=np.array(hi), box=box, **hits, **chunks)
    ao.close()
    ref.map_free(m)


if __name__ == "__main__":
    main()
