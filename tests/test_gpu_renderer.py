"""GPU: the host-side B200Renderer (openspades_b200/host/B200Renderer.*, SWRenderer's surface) driven the way
Client::DrawScene / HitTestDebugger drive SWRenderer (SetGameMap, StartScene, EndScene, FrameDone, ReadBitmap,
Flip; edits through the map's Set() -> GameMapChanged), frame for frame against the unmodified reference
SWRenderer where oracle/_ref travelled, else against the oracle restatement."""
import numpy as np
import pytest

import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import B200Renderer, VoxelMap

pytestmark = pytest.mark.gpu
FOG = (0.1, 0.1, 0.1)


def test_state_machine_errors_match_swrenderer():
    """Same misuse errors as SWRenderer's Ensure* guards (SWRenderer.cpp:520-547) and SetFramebuffer (:94-108)."""
    with pytest.raises(osb.OsbError, match="multiple of 8"):
        B200Renderer(100, 60)
    r = B200Renderer(64, 64, numerics=osb.NUM_IEEE)
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    sc = workloads.title_scene(64, 64, 68.0)
    with pytest.raises(osb.OsbError, match="not initialized"):
        r.set_game_map(m)
    with pytest.raises(osb.OsbError, match="not initialized"):
        r.start_scene(sc)
    r.init()
    assert r.screen() == (64.0, 64.0)
    with pytest.raises(osb.OsbError, match="outside of StartScene"):
        r.end_scene()
    with pytest.raises(osb.OsbError, match="outside of StartScene"):
        r.add_sprite()
    r.set_game_map(m)
    r.start_scene(sc)
    r.add_sprite()  # accepted, not drawn
    with pytest.raises(osb.OsbError, match="between StartScene"):
        r.start_scene(sc)
    with pytest.raises(osb.OsbError, match="between StartScene"):
        r.frame_done()
    with pytest.raises(osb.OsbError, match="between StartScene"):
        r.read_bitmap()
    r.end_scene()
    r.frame_done()
    r.flip()
    assert r.swap_count == 1
    r.shutdown()
    with pytest.raises(osb.OsbError, match="not valid"):
        r.frame_done()
    r.close()


def test_frames_with_edits_match_reference():
    w, h = 320, 240
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    have_ref = oracles.ref_available("sse2")
    port = oracles.Port()
    if have_ref:
        ref = oracles.Ref("sse2")
        rm = ref.map_from_arrays(m.solid.copy(), m.color.copy())
        rr = oracles.RefRenderer(ref, w, h, -1, 4)
        rr.set_map(rm)
        rr.set_fog(FOG, 128.0)
    r = B200Renderer(w, h, numerics=osb.NUM_SSE2, sw_threads=4)
    r.init()
    r.set_fog(FOG, 128.0)
    # before a map is set: clear colour, fog over the zero depth buffer leaves it untouched
    r.start_scene(workloads.title_scene(w, h, 68.0))
    r.end_scene()
    assert np.all((r.framebuffer() & 0xFFFFFF) == 0x7F7F7F)
    r.set_game_map(m)
    rng = np.random.default_rng(17)
    scenes = [workloads.title_scene(w, h, 68.0)] + list(workloads.random_scenes(5, m.solid, w, h, 68.0, seed=5))
    for fi, sc in enumerate(scenes):
        # edits in view of the camera between frames, through the map's Set() (listeners fire)
        ex, ey = int(sc["origin"][0]), int(sc["origin"][1])
        n_edit = 0 if fi == 0 else 60
        for _ in range(n_edit):
            x, y = (ex + int(rng.integers(-20, 21))) & 511, (ey + int(rng.integers(-20, 21))) & 511
            col = int(m.solid[x, y])
            top = (col & -col).bit_length() - 1
            eye_z = int(np.floor(sc["origin"][2]))
            if rng.random() < 0.5 and top < 62:
                z, s, c = top, False, 0
            else:
                z, s, c = max(top - 1, 0), True, 0x64000000 | int(rng.integers(0, 1 << 24))
                if x == ex and y == ey and z <= eye_z:
                    continue  # do not bury the camera
            m.set(x, y, z, s, c)
            if have_ref:
                ref.map_set(rm, x, y, z, s, c)
        r.start_scene(sc)
        r.end_scene()
        r.frame_done()
        assert (r.last_flushed > 0) == (n_edit > 0)
        fb, dp = r.framebuffer(), r.read_depth()
        if have_ref:
            wfb, wdp = rr.render(sc)
        else:
            wfb, wdp = port.render(m.solid.copy(), m.color.copy(), sc, w, h, FOG, 128.0, 2, 4, 1)
        same = ((fb & 0xFFFFFF) == (wfb & 0xFFFFFF)).mean()
        dsame = (dp.view(np.uint32) == wdp.view(np.uint32)).mean()
        assert same == 1.0 and dsame == 1.0, (fi, same, dsame)
        # ReadBitmap: RGBA, bottom-up, alpha 0xff (SWRenderer.cpp:1078-1098)
        bm = r.read_bitmap()
        exp = 0xFF000000 | (fb & 0xFF00) | ((fb & 0xFF) << 16) | ((fb & 0xFF0000) >> 16)
        assert np.array_equal(bm, exp[::-1])
        r.flip()
    # CastRay2Batch through the host map on the renderer's context sees the edited map
    v0, d = workloads.random_rays(50000, seed=3)
    got = r.castray2(m, v0, d, 256)
    want = port.castray2(m.solid.copy(), v0, d, 256, threads=8)
    assert got.tobytes() == want.tobytes()
    # radiosity leg with the reference's 8-chunk budget, then drain
    assert r.update_radiosity(8) == 8
    assert r.update_radiosity(-1) == 4096 - 8
    assert r.update_radiosity(8) == 0
    # GameMap::CastRay batch + the ambient-occlusion leg (GLAmbientShadowRenderer) on the same context
    d1 = d / np.linalg.norm(d, axis=1, keepdims=True)
    assert np.array_equal(r.castray(m, v0, d1, 64.0), port.castray(m.solid.copy(), v0, d1, 64.0, threads=8))
    assert r.update_ambient(8) == 8
    assert r.update_ambient(-1) == 4096 - 8
    assert r.update_ambient(8) == 0
    r.set_game_map(None)
    r.close()
    if have_ref:
        rr.set_map(None)
        rr.close()
        ref.map_free(rm)


def test_renderer_draws_models_and_lights_like_the_reference():
    """IRenderer surface: CreateModel / RenderModel / AddLight between StartScene and EndScene; the queued models and
    lights are drawn in submission order and the lists are cleared by EndScene (SWRenderer.cpp:565-576,709-745)."""
    import model_scenes as ms
    if not oracles.ref_available():
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("sse2")
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    w, h = 256, 160
    rm = ref.map_from_arrays(m.solid.copy(), m.color.copy())
    rr = oracles.RefRenderer(ref, w, h, -1, 4)
    rr.set_map(rm)
    rr.set_fog(FOG, 128.0)
    vm = oracles.RefVoxelModel(ref, kv6=open(ms.KV6[0], "rb").read())
    rmod = rr.create_model(vm)
    r = B200Renderer(w, h, numerics=osb.NUM_SSE2, sw_threads=4)
    r.init()
    r.set_fog(FOG, 128.0)
    r.set_game_map(m)
    s, c = vm.arrays()
    gmod = r.create_model(vm.w, vm.hgt, vm.d, vm.origin, s, c)
    sc = workloads.title_scene(w, h, 68.0)
    with pytest.raises(osb.OsbError, match="outside of StartScene"):
        r.add_light((0, 0, 0), 1.0, (1, 1, 1))
    pl = ms.placements(sc, 1, 3)
    par = np.zeros(len(pl), oracles.MODEL_PARAM_DTYPE)
    lrows = ms.lights(sc, 3)
    lg = np.zeros(len(lrows), oracles.LIGHT_DTYPE)
    r.start_scene(sc)
    for i, (_, mat, cc) in enumerate(pl):
        r.render_model(gmod, mat, cc)
        par["matrix"][i], par["customColor"][i] = mat, cc
    for i, (o, rad, col) in enumerate(lrows):
        r.add_light(o, rad, col)
        lg["origin"][i], lg["radius"][i], lg["color"][i] = o, rad, col
    r.add_light(sc["origin"], 50.0, (1, 1, 1), light_type=1)  # spotlight: ignored like the reference
    r.end_scene()
    r.frame_done()
    wfb, wdp = rr.render_scene(sc, [rmod] * len(pl), par, lg)
    assert np.array_equal(r.framebuffer() & 0xFFFFFF, wfb & 0xFFFFFF)
    assert np.array_equal(r.read_depth().view(np.uint32), wdp.view(np.uint32))
    # the next scene starts with empty lists
    r.start_scene(sc)
    r.end_scene()
    wfb2, _ = rr.render_scene(sc)
    assert np.array_equal(r.framebuffer() & 0xFFFFFF, wfb2 & 0xFFFFFF)
    assert not np.array_equal(wfb2 & 0xFFFFFF, wfb & 0xFFFFFF)
    # a model destroyed between RenderModel and EndScene is still drawn: the queue holds a reference until the scene
    # has been rendered (Handle<SWModel> in SWRenderer::Model, SWRenderer.h:87-91)
    r.start_scene(sc)
    for _, mat, cc in pl:
        r.render_model(gmod, mat, cc)
    for o, rad, col in lrows:
        r.add_light(o, rad, col)
    r.free_model(gmod)
    r.end_scene()
    assert np.array_equal(r.framebuffer() & 0xFFFFFF, wfb & 0xFFFFFF)
    # ... and a model may outlive its renderer
    late = r.create_model(vm.w, vm.hgt, vm.d, vm.origin, s, c)
    r.set_game_map(None)
    r.close()
    r.free_model(late)
    rr.free_model(rmod)
    vm.close()
    rr.set_map(None)
    rr.close()
    ref.map_free(rm)


def test_renderer_loads_and_saves_the_map_on_the_device():
    """B200Renderer::LoadGameMap (GameMap::Load on the device + SetGameMap) gives the frames of the host-decoded map,
    the returned VoxelMap takes edits like any other, and SaveGameMap is GameMap::Save of the edited map."""
    data = open(oracles.TITLE_VXL, "rb").read()
    port = oracles.Port()
    w, h = 128, 96
    sc = workloads.title_scene(w, h, 68.0)
    frames = []
    for device_load in (False, True):
        r = B200Renderer(w, h, numerics=osb.NUM_SSE2)
        r.init()
        r.set_fog(FOG, 128.0)
        if device_load:
            m = r.load_game_map(data)
        else:
            m = VoxelMap.load(data)
            r.set_game_map(m)
        out = []
        for step in range(2):
            r.start_scene(sc)
            r.end_scene()
            r.frame_done()
            out.append(r.read_bitmap().copy())
            for z in range(5, 31):  # a tower dead ahead; GameMap::Set -> GameMapChanged -> dirty column, visible next frame
                m.set(256, 266, z, True, 0x64112233 + z)
        saved = r.save_game_map()
        assert saved == port.vxl_save(m.solid.copy(), m.color.copy())
        frames.append(out)
        with pytest.raises(osb.OsbError):
            r.load_game_map(data[:1000])
        r.set_game_map(None)
        r.shutdown()
        r.close()
    for a, b in zip(*frames):
        assert np.array_equal(a, b)
    assert not np.array_equal(frames[0][0], frames[0][1])  # the edit is on screen


def test_voxel_map_wrapper_reports_falling_blocks_like_game_map_wrapper():
    """host/VoxelMapWrapper (client::GameMapWrapper's surface) on the renderer's context: a tower with an arm is built with
    AddBlock, cut with RemoveBlocks, and the reported blocks are those of the C restatement; the map edits reach the
    frame through the renderer's GameMapChanged listener."""
    from openspades_b200.hostlib import VoxelMapWrapper
    import links_cases
    port = oracles.Port()
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    ps, pc = m.solid.copy(), m.color.copy()
    pw = port.links_create(ps)
    r = B200Renderer(128, 96, numerics=osb.NUM_IEEE)
    r.init()
    r.set_fog(FOG, 128.0)
    r.set_game_map(m)
    w = VoxelMapWrapper(m, r)
    w.rebuild()
    x, y = 258, 300  # on the slope ahead of the title camera: the tower is on screen
    z0 = links_cases.top_of(ps, x, y)
    tower = [(x, y, z0 - 1 - k) for k in range(10)]
    arm = [(x + k, y, z0 - 10) for k in range(1, 5)]
    for c in tower + arm:
        w.add_block(*c, 0x64aa5511)
        port.links_add_block(pw, ps, pc, *c, 0x64aa5511)
    assert np.array_equal(m.solid, ps)
    sc = workloads.title_scene(128, 96, 68.0)
    r.start_scene(sc)
    r.end_scene()
    r.frame_done()
    before = r.read_bitmap().copy()
    got = w.remove_blocks([tower[3]])
    want = port.links_remove_blocks(pw, ps, [tower[3]])
    want = want[np.lexsort((want[:, 2], want[:, 1], want[:, 0]))]
    assert np.array_equal(got, want) and len(got) == 6 + 4
    for (cx, cy, cz) in got:  # World::ApplyBlockActions: the floating blocks are cleared
        m.set(int(cx), int(cy), int(cz), False, 0)
    r.start_scene(sc)
    r.end_scene()
    r.frame_done()
    assert not np.array_equal(before, r.read_bitmap())
    assert len(w.remove_blocks([tower[0]])) == 2  # what is left above the foot of the tower
    w.close()
    port.links_free(pw)
    r.set_game_map(None)
    r.shutdown()
    r.close()
