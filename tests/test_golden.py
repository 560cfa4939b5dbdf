"""Golden fixtures (tests/golden/*.npz, generated from the unmodified reference by tests/golden/make_golden.py).
CPU half: the oracle port reproduces them bit for bit (so the port stays pinned where oracle/_ref is absent).
GPU half (-m gpu): the CUDA path through the C ABI reproduces them bit for bit."""
import os

import numpy as np
import pytest

import oracles

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_arrays(port):
    return port.vxl_load(open(oracles.TITLE_VXL, "rb").read())


def _cases():
    g = np.load(os.path.join(GOLD, "castray2_title.npz"))
    r = np.load(os.path.join(GOLD, "radiosity_title.npz"))
    t = np.load(os.path.join(GOLD, "terrain_ieee_title.npz"))
    return g, r, t


def test_port_matches_golden(port, title_arrays):
    solid, color = title_arrays
    g, r, t = _cases()
    for s in (256, 64, 12):
        got = port.castray2(solid, g["v0"], g["dir"], s)
        assert np.array_equal(got.view(np.uint8).reshape(-1, 40), g[f"hits_{s}"]), s
    bm = port.shadow_bitmap(solid, color)
    assert np.array_equal(bm, r["shadow_bitmap"])
    f32, _ = port.radiosity_box(bm, tuple(r["box_lo"]), tuple(r["box_hi"]), 1, want_enc=False)
    assert np.array_equal(f32.view(np.uint32), r["box_f32"].view(np.uint32))
    for key, mode in (("chunk_16_16_2", 1), ("chunk_0_31_0", 1), ("chunk_mode2_16_16_2", 2)):
        cx, cy, cz = [int(v) for v in key.split("_")[-3:]]
        _, enc = port.radiosity_box(bm, (cx * 16, cy * 16, cz * 16), (cx * 16 + 15, cy * 16 + 15, cz * 16 + 15), mode,
                                    want_f32=False)
        assert np.array_equal(enc, r[key]), key
    fog = t["fog"]
    for i, sc in enumerate(t["scenes"]):
        fb, dp = port.render(solid, color, sc, 256, 256, fog[:3], float(fog[3]), 0, 4, 1)
        assert np.array_equal(fb & 0xFFFFFF, t["bgr"][i]) and np.array_equal(dp.view(np.uint32), t["depth"][i].view(np.uint32))


def _ambient_case():
    return np.load(os.path.join(GOLD, "ambient_title.npz"))


def test_port_matches_golden_ambient(port, title_arrays):
    """GameMap::CastRay + GLAmbientShadowRenderer fixtures (make_golden_ambient.py)."""
    solid, _ = title_arrays
    a = _ambient_case()
    for L in (16.0, 128.0, 3.5):
        assert np.array_equal(port.castray(solid, a["v0"], a["dir"], L), a[f"hits_{int(L * 10)}"]), L
    got = port.ao_evaluate_box(solid, a["rays"], tuple(int(v) for v in a["box_lo"]), tuple(int(v) for v in a["box_hi"]))
    assert np.array_equal(got.view(np.uint32), a["box"].view(np.uint32))
    st = port.ao_state_init()
    for key in ("chunk_16_16_2", "chunk_0_31_0", "chunk_31_0_3"):
        c = [int(v) for v in key.split("_")[1:]]
        assert np.array_equal(port.ao_update_chunk(solid, a["rays"], st, *c).view(np.uint32), a[key].view(np.uint32)), key


@pytest.mark.gpu
def test_cuda_matches_golden_ambient(title_arrays):
    import openspades_b200 as osb
    solid, color = title_arrays
    a = _ambient_case()
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        for L in (16.0, 128.0, 3.5):
            assert np.array_equal(ctx.castray(a["v0"], a["dir"], L), a[f"hits_{int(L * 10)}"]), L
        ctx.ambient_set_rays(a["rays"])
        got = ctx.ambient_evaluate_box(tuple(int(v) for v in a["box_lo"]), tuple(int(v) for v in a["box_hi"]))
        assert np.array_equal(got.view(np.uint32), a["box"].view(np.uint32))
        ctx.ambient_update(-1)
        for key in ("chunk_16_16_2", "chunk_0_31_0", "chunk_31_0_3"):
            c = [int(v) for v in key.split("_")[1:]]
            assert np.array_equal(ctx.ambient_read_chunk(*c).view(np.uint32), a[key].view(np.uint32)), key


@pytest.mark.gpu
def test_cuda_matches_golden(title_arrays):
    import openspades_b200 as osb
    solid, color = title_arrays
    g, r, t = _cases()
    with osb.Context(0, osb.NUM_IEEE, sw_threads=4, radiosity_mode=1) as ctx:
        ctx.map_upload(solid, color)
        for s in (256, 64, 12):
            got = ctx.castray2(g["v0"], g["dir"], s)
            assert np.array_equal(got.view(np.uint8).reshape(-1, 40), g[f"hits_{s}"]), s
        assert np.array_equal(ctx.shadow_bitmap(), r["shadow_bitmap"])
        got = ctx.radiosity_evaluate_box(tuple(int(v) for v in r["box_lo"]), tuple(int(v) for v in r["box_hi"]))
        assert np.array_equal(got.view(np.uint32), r["box_f32"].view(np.uint32))
        ctx.radiosity_update(-1)
        for key in ("chunk_16_16_2", "chunk_0_31_0"):
            cx, cy, cz = [int(v) for v in key.split("_")[-3:]]
            assert np.array_equal(ctx.radiosity_read_chunk(cx, cy, cz), r[key]), key
        fog = t["fog"]
        fb, dp = ctx.render(t["scenes"].view(osb.SCENE_DTYPE).reshape(-1), 256, 256, fog[:3], float(fog[3]))
        assert np.array_equal(fb & 0xFFFFFF, t["bgr"]) and np.array_equal(dp.view(np.uint32), t["depth"].view(np.uint32))
    with osb.Context(0, osb.NUM_IEEE, radiosity_mode=2) as ctx:
        ctx.map_upload(solid, color)
        ctx.radiosity_update(-1)
        assert np.array_equal(ctx.radiosity_read_chunk(16, 16, 2), r["chunk_mode2_16_16_2"])


@pytest.mark.gpu
def test_cuda_matches_golden_models(title_arrays):
    """Frames with voxel models + dynamic lights drawn by the unmodified SWRenderer (make_golden_models.py, V-IEEE)."""
    import openspades_b200 as osb
    solid, color = title_arrays
    g = np.load(os.path.join(GOLD, "models_title.npz"))
    fog = g["fog"]
    with osb.Context(0, osb.NUM_IEEE, sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        mods = []
        for i in range(int(g["n_models"])):
            w, h, d = (int(v) for v in g[f"m{i}_dims"])
            mo = ctx.model_create(w, h, d, g[f"m{i}_origin"], g[f"m{i}_solid"], g[f"m{i}_colors"])
            data, addr = ctx.model_render_data(mo, w * h)
            assert np.array_equal(data, g[f"m{i}_data"]) and np.array_equal(addr, g[f"m{i}_addr"])
            mods.append(mo)
        H, W = g["s0_bgr"].shape
        for si, sc in enumerate(g["scenes"].view(osb.SCENE_DTYPE).reshape(-1)):
            idx = g[f"s{si}_model_index"]
            inst = np.zeros(len(idx), osb.MODEL_INSTANCE_DTYPE)
            inst["model"] = [mods[int(k)] for k in idx]
            inst["matrix"] = g[f"s{si}_matrix"]
            inst["customColor"] = g[f"s{si}_custom"]
            lg = np.ascontiguousarray(g[f"s{si}_lights"]).view(osb.LIGHT_DTYPE).reshape(-1)
            fb, dp = ctx.render_scene(sc, W, H, fog[:3], float(fog[3]), inst, lg)
            assert np.array_equal(fb & 0xFFFFFF, g[f"s{si}_bgr"]), si
            assert np.array_equal(dp.view(np.uint32), g[f"s{si}_depth"].view(np.uint32)), si
        for mo in mods:
            ctx.model_destroy(mo)
