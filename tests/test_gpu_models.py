"""GPU parity of the full EndScene -- terrain, voxel models, dynamic lights, fog -- (osb200_model_create /
osb200_render_scene through the C ABI) against the unmodified reference SWRenderer (oracle/_ref: SWModel,
SWModelRenderer, SWRenderer::AddLight / ApplyDynamicLight / ApplyFog) and the committed fixture.
Bar: depth buffer (= which splat won every pixel) bit-exact, colours within +-1 LSB; written to be bit-identical."""
import os

import numpy as np
import pytest

import model_scenes as ms
import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu
FOG = (0.8, 1.0, 1.0)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def title_map():
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    return m.solid.copy(), m.color.copy()


def _ref_models(ref):
    vms = [oracles.RefVoxelModel(ref, kv6=open(p, "rb").read()) for p in ms.KV6]
    vms.append(oracles.RefVoxelModel(ref, arrays=ms.synthetic_model(5)))
    vms.append(oracles.RefVoxelModel(ref, arrays=ms.synthetic_model(6, 30, 30, 40)))
    return vms


def _case(scene, n_models, seed):
    pl = ms.placements(scene, n_models, seed)
    inst = np.zeros(len(pl), osb.MODEL_INSTANCE_DTYPE)
    par = np.zeros(len(pl), oracles.MODEL_PARAM_DTYPE)
    for i, (mi, m, cc) in enumerate(pl):
        inst["matrix"][i] = par["matrix"][i] = m
        inst["customColor"][i] = par["customColor"][i] = cc
    lg = np.zeros(len(ms.lights(scene, seed)), osb.LIGHT_DTYPE)
    for i, (o, r, c) in enumerate(ms.lights(scene, seed)):
        lg["origin"][i], lg["radius"][i], lg["color"][i] = o, r, c
    return pl, inst, par, lg


@pytest.mark.parametrize("variant,numerics", [("sse2", osb.NUM_SSE2), ("ieee", osb.NUM_IEEE)])
def test_scene_with_models_and_lights_matches_reference(title_map, variant, numerics):
    if not oracles.ref_available(variant):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref(variant)
    solid, color = title_map
    rm = ref.map_from_arrays(solid, color)
    w, h = 320, 200
    rr = oracles.RefRenderer(ref, w, h, -1, 4)
    rr.set_map(rm)
    rr.set_fog(FOG, 128.0)
    vms = _ref_models(ref)
    rmods = [rr.create_model(v) for v in vms]
    scenes = [workloads.title_scene(w, h, 68.0)] + list(workloads.random_scenes(5, solid, w, h, 68.0, seed=11))
    with osb.Context(0, numerics, sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        gm = []
        for v, rmod in zip(vms, rmods):
            s, c = v.arrays()
            g = ctx.model_create(v.w, v.hgt, v.d, v.origin, s, c)
            gm.append(g)
            gd, ga = ctx.model_render_data(g, v.w * v.hgt)
            wd, wa = rr.model_render_data(rmod, v.w * v.hgt)
            assert np.array_equal(gd, wd) and np.array_equal(ga, wa)  # SWModel ctor
        for si, sc in enumerate(scenes):
            pl, inst, par, lg = _case(sc, len(vms), 100 + si)
            inst["model"] = [gm[mi] for mi, _, _ in pl]
            for what in ("both", "models", "lights", "none"):
                ii = inst if what in ("both", "models") else None
                ll = lg if what in ("both", "lights") else None
                fb, dp = ctx.render_scene(sc, w, h, FOG, 128.0, ii, ll)
                wfb, wdp = rr.render_scene(sc, [rmods[mi] for mi, _, _ in pl] if ii is not None else [],
                                           par if ii is not None else None,
                                           lg.view(oracles.LIGHT_DTYPE) if ll is not None else None)
                dsame = (dp.view(np.uint32) == wdp.view(np.uint32)).mean()
                d = np.abs(((fb[..., None] >> np.array([0, 8, 16])) & 0xFF).astype(int)
                           - ((wfb[..., None] >> np.array([0, 8, 16])) & 0xFF).astype(int))
                same = ((fb & 0xFFFFFF) == (wfb & 0xFFFFFF)).mean()
                assert dsame == 1.0 and d.max() <= 1 and same == 1.0, (variant, si, what, dsame, same, d.max())
            if si == 0:
                fb0, _ = ctx.render_scene(sc, w, h, FOG, 128.0)
                fb1, _ = ctx.render_scene(sc, w, h, FOG, 128.0, inst, lg)
                assert ((fb0 & 0xFFFFFF) != (fb1 & 0xFFFFFF)).mean() > 0.05  # the case really draws something
        for g in gm:
            ctx.model_destroy(g)
    for m in rmods:
        rr.free_model(m)
    for v in vms:
        v.close()
    rr.set_map(None)
    rr.close()
    ref.map_free(rm)


def test_render_scene_argument_errors(title_map):
    solid, color = title_map
    with osb.Context(0, osb.NUM_IEEE) as ctx, osb.Context(0, osb.NUM_IEEE) as other:
        ctx.map_upload(solid, color)
        sc = workloads.title_scene(64, 64, 68.0)
        with pytest.raises(osb.OsbError, match="multiple of 8"):
            ctx.render_scene(sc, 60, 64, FOG, 128.0)
        w, h, d, o, s, c = ms.synthetic_model(1)
        with pytest.raises(osb.OsbError, match="depth"):
            ctx.model_create(4, 4, 65, o, np.zeros((4, 4), np.uint64), np.zeros((4, 4, 65), np.uint32))
        foreign = other.model_create(w, h, d, o, s, c)
        inst = np.zeros(1, osb.MODEL_INSTANCE_DTYPE)
        inst["model"] = foreign
        inst["matrix"][0] = ms.matrix((0, 0, 0), (1, 1, 1))
        with pytest.raises(osb.OsbError, match="model of this context"):
            ctx.render_scene(sc, 64, 64, FOG, 128.0, inst)
        other.model_destroy(foreign)
        fb, dp = ctx.render_scene(sc, 64, 64, FOG, 128.0, np.zeros(0, osb.MODEL_INSTANCE_DTYPE))  # empty lists
        fb2, dp2 = ctx.render(sc, 64, 64, FOG, 128.0)
        assert np.array_equal(fb & 0xFFFFFF, fb2[0] & 0xFFFFFF) and np.array_equal(dp, dp2[0])


def test_bench_game_frame_1080p_matches_port(title_map):
    """The bench's in-game frame (1080p, 64 instances of 24x24x40 models, 8 lights; bench.py extra frame_1080p_game) and a
    denser one (256 instances, 32 lights) against the C restatement, bit for bit -- the full-size case of this path."""
    solid, color = title_map
    port = oracles.Port()
    w, h = 1920, 1080
    sc = workloads.title_scene(w, h, 68.0)
    mdls = [workloads.blob_model(11 + k) for k in range(3)]
    pms = [port.model_create(*m) for m in mdls]
    with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        gms = [ctx.model_create(*m) for m in mdls]
        for n_inst, n_lights in ((64, 8), (256, 32)):
            inst, lights = workloads.game_frame(sc, gms, n_inst, n_lights)
            fb, dp = ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights)
            par = np.zeros(n_inst, oracles.MODEL_PARAM_DTYPE)
            par["matrix"], par["customColor"] = inst["matrix"], inst["customColor"]
            wfb, wdp = port.render_scene(solid, color, sc, w, h, (0.1, 0.1, 0.1), 128.0, 2, [pms[i % 3] for i in range(n_inst)],
                                         par, lights.view(oracles.LIGHT_DTYPE), threads=4)
            assert np.array_equal(dp.view(np.uint32), wdp.view(np.uint32)), n_inst
            assert np.array_equal(fb & 0xFFFFFF, wfb & 0xFFFFFF), n_inst
            terrain_dp = port.render(solid, color, sc, w, h, (0.1, 0.1, 0.1), 128.0, 2, 4, 1)[1]
            assert (wdp.view(np.uint32) != terrain_dp.view(np.uint32)).sum() > 10000  # the models are on screen
        for g in gms:
            ctx.model_destroy(g)
    for p in pms:
        port.model_free(p)
