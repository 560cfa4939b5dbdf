"""Pins oracle/orc_models.c (CPU restatement of SWModel / SWModelRenderer / AddLight / ApplyDynamicLight inside
EndScene) against the unmodified reference compiled in place (oracle/_ref) and against the committed fixture
tests/golden/models_title.npz.  Bar: every word bit-identical."""
import os

import numpy as np
import pytest

import model_scenes as ms
import oracles
from openspades_b200 import workloads

FOG = (0.8, 1.0, 1.0)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NUMERICS = {"sse2": 2, "ieee": 0}


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_arrays(port):
    return port.vxl_load(open(oracles.TITLE_VXL, "rb").read())


def _case(scene, n_models, seed):
    pl = ms.placements(scene, n_models, seed)
    par = np.zeros(len(pl), oracles.MODEL_PARAM_DTYPE)
    for i, (_, m, cc) in enumerate(pl):
        par["matrix"][i], par["customColor"][i] = m, cc
    rows = ms.lights(scene, seed)
    lg = np.zeros(len(rows), oracles.LIGHT_DTYPE)
    for i, (o, r, c) in enumerate(rows):
        lg["origin"][i], lg["radius"][i], lg["color"][i] = o, r, c
    return pl, par, lg


@pytest.mark.parametrize("variant", ["sse2", "ieee"])
def test_port_scene_with_models_and_lights_matches_reference(port, title_arrays, variant):
    if not oracles.ref_available(variant):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref(variant)
    solid, color = title_arrays
    rm = ref.map_from_arrays(solid, color)
    w, h = 160, 104
    rr = oracles.RefRenderer(ref, w, h, -1, 4)
    rr.set_map(rm)
    rr.set_fog(FOG, 128.0)
    vms = [oracles.RefVoxelModel(ref, kv6=open(p, "rb").read()) for p in ms.KV6]
    vms.append(oracles.RefVoxelModel(ref, arrays=ms.synthetic_model(5)))
    rmods = [rr.create_model(v) for v in vms]
    pmods = []
    for v, rmod in zip(vms, rmods):
        s, c = v.arrays()
        pm = port.model_create(v.w, v.hgt, v.d, v.origin, s, c)
        pmods.append(pm)
        pd, pa = port.model_render_data(pm, v.w * v.hgt)
        wd, wa = rr.model_render_data(rmod, v.w * v.hgt)
        assert np.array_equal(pd, wd) and np.array_equal(pa, wa)  # SWModel ctor
    scenes = [workloads.title_scene(w, h, 68.0)] + list(workloads.random_scenes(2, solid, w, h, 68.0, seed=11))
    depth = np.zeros((h, w), np.float32)  # SWRenderer::depthBuffer persists from frame to frame
    for si, sc in enumerate(scenes):
        pl, par, lg = _case(sc, len(vms), 100 + si)
        for what in ("both", "models", "lights"):
            use_m, use_l = what in ("both", "models"), what in ("both", "lights")
            wfb, wdp = rr.render_scene(sc, [rmods[mi] for mi, _, _ in pl] if use_m else [], par if use_m else None,
                                       lg if use_l else None)
            fb, dp = port.render_scene(solid, color, sc, w, h, FOG, 128.0, NUMERICS[variant],
                                       [pmods[mi] for mi, _, _ in pl] if use_m else [], par if use_m else None,
                                       lg if use_l else None, threads=4, depth=depth)
            assert np.array_equal(dp.view(np.uint32), wdp.view(np.uint32)), (variant, si, what)
            assert np.array_equal(fb & 0xFFFFFF, wfb & 0xFFFFFF), (variant, si, what)
    for pm in pmods:
        port.model_free(pm)
    for rmod in rmods:
        rr.free_model(rmod)
    for v in vms:
        v.close()
    rr.set_map(None)
    rr.close()
    ref.map_free(rm)


def test_port_matches_golden_models(port, title_arrays):
    """The committed frames of the unmodified SWRenderer (make_golden_models.py, V-IEEE, r_swNumThreads 4)."""
    solid, color = title_arrays
    g = np.load(os.path.join(GOLD, "models_title.npz"))
    fog = g["fog"]
    mods = []
    for i in range(int(g["n_models"])):
        w, h, d = (int(v) for v in g[f"m{i}_dims"])
        mo = port.model_create(w, h, d, g[f"m{i}_origin"], g[f"m{i}_solid"], g[f"m{i}_colors"])
        data, addr = port.model_render_data(mo, w * h)
        assert np.array_equal(data, g[f"m{i}_data"]) and np.array_equal(addr, g[f"m{i}_addr"])
        mods.append(mo)
    H, W = g["s0_bgr"].shape
    depth = np.zeros((H, W), np.float32)
    for si, sc in enumerate(g["scenes"].view(oracles.SCENE_DTYPE).reshape(-1)):
        idx = g[f"s{si}_model_index"]
        par = np.zeros(len(idx), oracles.MODEL_PARAM_DTYPE)
        par["matrix"], par["customColor"] = g[f"s{si}_matrix"], g[f"s{si}_custom"]
        lg = np.ascontiguousarray(g[f"s{si}_lights"]).view(oracles.LIGHT_DTYPE).reshape(-1)
        fb, dp = port.render_scene(solid, color, sc, W, H, fog[:3], float(fog[3]), 0, [mods[int(k)] for k in idx], par, lg,
                                   threads=4, depth=depth)
        assert np.array_equal(fb & 0xFFFFFF, g[f"s{si}_bgr"]), si
        assert np.array_equal(dp.view(np.uint32), g[f"s{si}_depth"].view(np.uint32)), si
    for mo in mods:
        port.model_free(mo)


def test_port_without_map_draws_models_over_the_clear_colour(port):
    """SetGameMap(null): EndScene skips the terrain (SWRenderer.cpp:720) and models land on the 0x7f7f7f clear."""
    w, h = 64, 48
    sc = workloads.title_scene(w, h, 68.0)
    mw, mh, md, origin, s, c = ms.synthetic_model(3)
    mo = port.model_create(mw, mh, md, origin, s, c)
    par = np.zeros(1, oracles.MODEL_PARAM_DTYPE)
    pos = np.asarray(sc["origin"]) + np.asarray(sc["axis"][2]) * 6.0
    par["matrix"][0] = ms.matrix(pos, (0.2, 0.2, 0.2))
    depth = np.full((h, w), 1e9, np.float32)
    fb, dp = port.render_scene(None, None, sc, w, h, FOG, 128.0, 0, [mo], par, None, depth=depth)
    assert (dp < 1e9).any() and (dp == 1e9).any()
    port.model_free(mo)
