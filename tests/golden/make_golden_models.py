#!/usr/bin/env python
"""Generates tests/golden/models_title.npz from the UNMODIFIED reference (oracle/_ref/liboref_ieee.so): frames of
Title.vxl with voxel models and dynamic point lights drawn by SWRenderer (SWModel, SWModelRenderer, AddLight,
ApplyDynamicLight, ApplyFog), V-IEEE build (CPU independent), r_swNumThreads = 4.  The models' VoxelModel arrays (KV6
files decoded by the reference's loader + two synthetic ones) and SWModel render data are stored with the frames."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import model_scenes as ms  # noqa: E402
import oracles  # noqa: E402
from openspades_b200 import workloads  # noqa: E402

W, H = 192, 128
FOG = (0.8, 1.0, 1.0)


def main():
    ref = oracles.Ref("ieee")
    m = ref.map_load_vxl(open(oracles.TITLE_VXL, "rb").read())
    solid, _ = ref.map_arrays(m)
    rr = oracles.RefRenderer(ref, W, H, -1, 4)
    rr.set_map(m)
    rr.set_fog(FOG, 128.0)
    vms = [oracles.RefVoxelModel(ref, kv6=open(p, "rb").read()) for p in ms.KV6]
    vms.append(oracles.RefVoxelModel(ref, arrays=ms.synthetic_model(5)))
    mods = [rr.create_model(v) for v in vms]
    out = {"n_models": np.array(len(vms)), "fog": np.array(FOG + (128.0,), np.float32)}
    for i, (v, mod) in enumerate(zip(vms, mods)):
        s, c = v.arrays()
        d, a = rr.model_render_data(mod, v.w * v.hgt)
        out.update({f"m{i}_dims": np.array([v.w, v.hgt, v.d]), f"m{i}_origin": v.origin, f"m{i}_solid": s, f"m{i}_colors": c,
                    f"m{i}_data": d, f"m{i}_addr": a})
    scenes = np.stack([workloads.title_scene(W, H, 68.0), workloads.random_scenes(3, solid, W, H, 68.0, seed=21)[2]])
    out["scenes"] = scenes
    for si, sc in enumerate(scenes):
        pl = ms.placements(sc, len(vms), 40 + si)
        par = np.zeros(len(pl), oracles.MODEL_PARAM_DTYPE)
        for i, (_, mat, cc) in enumerate(pl):
            par["matrix"][i], par["customColor"][i] = mat, cc
        rows = ms.lights(sc, 40 + si)
        lg = np.zeros(len(rows), oracles.LIGHT_DTYPE)
        for i, (o, r, c) in enumerate(rows):
            lg["origin"][i], lg["radius"][i], lg["color"][i] = o, r, c
        fb, dp = rr.render_scene(sc, [mods[mi] for mi, _, _ in pl], par, lg)
        out.update({f"s{si}_model_index": np.array([mi for mi, _, _ in pl]), f"s{si}_matrix": par["matrix"],
                    f"s{si}_custom": par["customColor"], f"s{si}_lights": lg.view(np.float32).reshape(len(lg), 7),
                    f"s{si}_bgr": fb & 0xFFFFFF, f"s{si}_depth": dp})
    np.savez_compressed(os.path.join(HERE, "models_title.npz"), **out)
    for mod in mods:
        rr.free_model(mod)
    for v in vms:
        v.close()
    rr.set_map(None)
    rr.close()
    ref.map_free(m)


if __name__ == "__main__":
    main()
