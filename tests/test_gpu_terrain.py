"""GPU parity of the terrain pass + fog (osb200_render through the C ABI) against the oracle
restatement (oracle/liboracle.so) and, where it travelled, the unmodified reference (oracle/_ref).

Bar (BASELINE.json north_star): >= 95 % of pixels bit-identical, the rest within +-1 LSB per channel;
depth ordering bit-exact.  The kernels are written to be bit-identical, so the tests assert 100 %."""
import os

import numpy as np
import pytest

import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu

FOG = (0.1, 0.1, 0.1)
NUMERICS = {"ieee": osb.NUM_IEEE, "sse": osb.NUM_SSE, "sse2": osb.NUM_SSE2}


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_map():
    return VoxelMap.load_file(oracles.TITLE_VXL)


@pytest.fixture(scope="module")
def synth():
    return workloads.synthetic_map(7)


def _report(tag, fb, dp, rfb, rdp):
    same = (fb & 0xFFFFFF) == (rfb & 0xFFFFFF)
    dsame = dp.view(np.uint32) == rdp.view(np.uint32)
    frac, dfrac = same.mean(), dsame.mean()
    if frac < 1 or dfrac < 1:
        d = np.abs(((fb[..., None] >> np.array([0, 8, 16], np.uint32)) & 0xFF).astype(int)
                   - ((rfb[..., None] >> np.array([0, 8, 16], np.uint32)) & 0xFF).astype(int)).max(axis=-1)
        print(f"{tag}: rgb identical {frac:.6f}, within1 {(d <= 1).mean():.6f}, max {d.max()}, depth identical {dfrac:.6f}")
    return frac, dfrac


@pytest.mark.parametrize("mode", ["ieee", "sse", "sse2"])
@pytest.mark.parametrize("size", [(800, 600, 30.0), (256, 256, 68.0), (1920, 1080, 68.0)])
def test_title_scene_matches_oracle(port, title_map, mode, size):
    w, h, fov = size
    solid, color = title_map.solid.copy(), title_map.color.copy()
    sc = workloads.title_scene(w, h, fov)
    with osb.Context(0, NUMERICS[mode], sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        fb, dp = ctx.render(sc, w, h, FOG, 128.0)
    rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, NUMERICS[mode], 4, 1)
    frac, dfrac = _report(f"title {mode} {w}x{h}", fb[0], dp[0], rfb, rdp)
    assert frac == 1.0 and dfrac == 1.0
    if mode == "sse2":  # full 32-bit word including the reference's junk X byte
        assert np.array_equal(fb[0], rfb)


@pytest.mark.parametrize("mode", ["ieee", "sse2"])
def test_random_poses_batch_matches_oracle(port, synth, mode):
    """40 seeded poses (pitch +-60 deg: both the yawRange = pi and the pole-visible 2 pi case), rendered as
    one batch (device-side recurrences) and compared view by view."""
    solid, color = synth
    w = h = 256
    scenes = workloads.random_scenes(40, solid, w, h, 68.0, seed=4321)
    with osb.Context(0, NUMERICS[mode], sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        fb, dp = ctx.render(scenes, w, h, (0.8, 1.0, 1.0), 128.0)
        # the same views one at a time (host-side recurrences) must give the same bits
        fb1, dp1 = ctx.render(scenes[:3], w, h, (0.8, 1.0, 1.0), 128.0)
    assert np.array_equal(fb[:3], fb1) and np.array_equal(dp[:3].view(np.uint32), dp1.view(np.uint32))
    worst = 1.0
    for i in range(len(scenes)):
        rfb, rdp = port.render(solid, color, scenes[i], w, h, (0.8, 1.0, 1.0), 128.0, NUMERICS[mode], 4, 1)
        frac, dfrac = _report(f"pose {i} {mode}", fb[i], dp[i], rfb, rdp)
        worst = min(worst, frac, dfrac)
    assert worst == 1.0


def test_against_unmodified_reference(title_map):
    """The real thing, when oracle/_ref travelled to this box: SWRenderer (SSE2 level, 4 threads)."""
    if not oracles.ref_available("sse2"):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("sse2")
    solid, color = title_map.solid.copy(), title_map.color.copy()
    m = ref.map_from_arrays(solid, color)
    with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:
        ctx.map_upload(solid, color)
        for (w, h, fov) in ((800, 600, 30.0), (1920, 1080, 68.0)):
            r = oracles.RefRenderer(ref, w, h, -1, 4)
            r.set_map(m)
            r.set_fog(FOG, 128.0)
            scenes = [workloads.title_scene(w, h, fov)] + list(workloads.random_scenes(4, solid, w, h, fov, seed=11))
            for i, sc in enumerate(scenes):
                rfb, rdp = r.render(sc)
                fb, dp = ctx.render(sc, w, h, FOG, 128.0)
                frac, dfrac = _report(f"ref {w}x{h} #{i}", fb[0], dp[0], rfb, rdp)
                assert frac == 1.0 and dfrac == 1.0
            r.set_map(None)
            r.close()
    ref.map_free(m)


def test_eye_inside_solid_and_bad_size(title_map):
    solid, color = title_map.solid.copy(), title_map.color.copy()
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        sc = workloads.title_scene(64, 64, 68.0)
        sc["origin"] = (256.5, 256.5, 62.5)  # inside the ground
        fb, dp = ctx.render(sc, 64, 64, FOG, 128.0)
        # cleared to 0x7f7f7f then fogged with the stale (zero) depth -> unchanged colour, depth untouched
        assert np.all((fb & 0xFFFFFF) == 0x7F7F7F) and np.all(dp == 0)
        with pytest.raises(osb.OsbError, match="multiple of 8"):
            ctx.render(sc, 100, 60, FOG, 128.0)


def test_small_launch_lane_dealing_does_not_change_the_frame(tmp_path):
    """Small launches are walked differently from batches (OSB200_LINE_MIN_LANES / OSB200_COOP, read once per process): fewer
    lines per warp task, or the cooperative walker k_walk_coop with a whole warp per yaw line.  The same views rendered in
    fresh processes under each policy give identical frames + depth."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np\n"
        f"sys.path.insert(0, {root!r})\n"
        "import openspades_b200 as osb\n"
        "from openspades_b200 import workloads\n"
        "from openspades_b200.hostlib import VoxelMap\n"
        "m = VoxelMap.load_file(workloads.TITLE_VXL)\n"
        "with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:\n"
        "    ctx.map_upload(m.solid.copy(), m.color.copy())\n"
        "    S = workloads.random_scenes(3, m.solid, 320, 200, seed=9)\n"
        "    fb, dp = ctx.render(S, 320, 200, (0.8, 1.0, 1.0), 128.0)\n"
        "    sc = workloads.title_scene(1920, 1080, 68.0)\n"
        "    fb2, dp2 = ctx.render(sc, 1920, 1080, (0.1, 0.1, 0.1), 128.0)\n"
        "np.savez(sys.argv[1], fb=fb, dp=dp, fb2=fb2, dp2=dp2)\n")
    outs = []
    # lane-per-line walker with 32 / 4 / 1 lines per warp task, then the cooperative walker (one warp per line), forced for
    # the 3-view batch too (by default it serves launches of fewer than 64 lines per SM)
    for lanes, coop in (("32", "0"), ("4", "0"), ("1", "0"), ("4", "1")):
        out = str(tmp_path / f"lanes{lanes}_{coop}.npz")
        env = dict(os.environ, OSB200_LINE_MIN_LANES=lanes, OSB200_COOP=coop)
        subprocess.run([sys.executable, "-c", code, out], check=True, env=env, timeout=300)
        outs.append(np.load(out))
    for other in outs[1:]:
        for k in ("fb", "dp", "fb2", "dp2"):
            assert np.array_equal(outs[0][k].view(np.uint32), other[k].view(np.uint32)), k


@pytest.mark.parametrize("threads,under", [(1, 1), (32, 1), (7, 2)])
def test_batch_with_device_side_tables(port, title_map, threads, under):
    """More than 8 views take the device-side recurrences (k_view_prep: one thread per (job, view), 2 + 2 T jobs) instead of
    the host loop the small tests above go through: the same r_swNumThreads / r_swUndersampling dependence, checked on a
    batch of 24 views against the oracle port."""
    solid, color = title_map.solid.copy(), title_map.color.copy()
    w = h = 256
    scenes = workloads.random_scenes(24, solid, w, h, 68.0, seed=77)
    with osb.Context(0, osb.NUM_SSE2, sw_threads=threads, undersampling=under) as ctx:
        ctx.map_upload(solid, color)
        fbs, dps = ctx.render(scenes, w, h, (0.8, 1.0, 1.0), 128.0)
    for i in (0, 7, 16, 23):
        pfb, pdp = port.render(solid, color, scenes[i], w, h, (0.8, 1.0, 1.0), 128.0, osb.NUM_SSE2, threads, under)
        assert _report(f"batch threads {threads} under {under} #{i}", fbs[i], dps[i], pfb, pdp) == (1.0, 1.0)


def test_view_pipeline_gives_the_same_frames(tmp_path, port, title_map):
    """k_view_pipeline (OSB200_PIPELINE = window in views, off by default: measured slower) walks and
    gathers a view back to back inside one persistent kernel.  A batch large enough to take that path, rendered in fresh
    processes with the pipeline off, with an unlimited window, with a window of 8 views (admission waits) and with roles
    per SM, gives identical frames + depth; a few of them are checked against the oracle port."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, numpy as np\n"
        f"sys.path.insert(0, {root!r})\n"
        "import openspades_b200 as osb\n"
        "from openspades_b200 import workloads\n"
        "from openspades_b200.hostlib import VoxelMap\n"
        "m = VoxelMap.load_file(workloads.TITLE_VXL)\n"
        "with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:\n"
        "    ctx.map_upload(m.solid.copy(), m.color.copy())\n"
        "    S = workloads.random_scenes(96, m.solid, 256, 256, seed=31)\n"
        "    S[5]['origin'] = (256.5, 256.5, 62.5)\n"  # a view with the eye inside the ground
        "    ctx.timing_enable(True)\n"
        "    fb, dp = ctx.render(S, 256, 256, (0.8, 1.0, 1.0), 128.0)\n"
        "    names = sorted(k for k, v in ctx.timing_read().items() if v[0])\n"
        "np.savez(sys.argv[1], fb=fb, dp=dp, names=np.array(names))\n")
    outs = []
    for window, split in (("0", "0"), ("4096", "0"), ("8", "0"), ("4096", "10")):
        out = str(tmp_path / f"pipe{window}_{split}.npz")
        env = dict(os.environ, OSB200_PIPELINE=window, OSB200_PIPE_SPLIT=split)
        subprocess.run([sys.executable, "-c", code, out], check=True, env=env, timeout=300)
        outs.append(np.load(out))
        assert ("k_view_pipeline" in list(outs[-1]["names"])) == (window != "0"), list(outs[-1]["names"])
    for other in outs[1:]:
        for k in ("fb", "dp"):
            assert np.array_equal(outs[0][k].view(np.uint32), other[k].view(np.uint32)), k
    solid, color = title_map.solid.copy(), title_map.color.copy()
    S = workloads.random_scenes(96, solid, 256, 256, seed=31)
    for i in (0, 40, 95):
        pfb, pdp = port.render(solid, color, S[i], 256, 256, (0.8, 1.0, 1.0), 128.0, osb.NUM_SSE2, 4, 1)
        assert _report(f"pipeline view {i} vs port", outs[1]["fb"][i], outs[1]["dp"][i], pfb, pdp) == (1.0, 1.0)


@pytest.mark.parametrize("under", [2, 4])
@pytest.mark.parametrize("size", [(800, 600, 30.0), (256, 256, 68.0)])
def test_undersampling_matches_reference(port, title_map, under, size):
    """r_swUndersampling 2 / 4 (SWMapRenderer.cpp:37,1110-1112: numLines /= under; store replication :889-911,987-1009):
    osb200_opts.undersampling against the unmodified SWRenderer with the cvar set (when oracle/_ref travelled) and the port."""
    w, h, fov = size
    solid, color = title_map.solid.copy(), title_map.color.copy()
    scenes = [workloads.title_scene(w, h, fov)] + list(workloads.random_scenes(3, solid, w, h, fov, seed=21))
    ref = oracles.Ref("sse2") if oracles.ref_available("sse2") else None
    with osb.Context(0, osb.NUM_SSE2, sw_threads=4, undersampling=under) as ctx:
        ctx.map_upload(solid, color)
        fbs, dps = ctx.render(np.stack(scenes), w, h, FOG, 128.0)
    if ref is not None:
        m = ref.map_from_arrays(solid, color)
        r = oracles.RefRenderer(ref, w, h, -1, 4, under)
        r.set_map(m)
        r.set_fog(FOG, 128.0)
    for i, sc in enumerate(scenes):
        pfb, pdp = port.render(solid, color, sc, w, h, FOG, 128.0, osb.NUM_SSE2, 4, under)
        assert _report(f"under {under} {w}x{h} #{i} vs port", fbs[i], dps[i], pfb, pdp) == (1.0, 1.0)
        if ref is not None:
            rfb, rdp = r.render(sc)
            assert _report(f"under {under} {w}x{h} #{i} vs ref", fbs[i], dps[i], rfb, rdp) == (1.0, 1.0)
    if ref is not None:
        r.set_map(None)
        r.close()
        oracles.RefRenderer(ref, 8, 8, -1, 4, 0).close()  # leave the process-wide cvar at its default
        ref.map_free(m)


@pytest.mark.parametrize("threads", [1, 8, 32])
def test_sw_threads_matches_reference(port, title_map, threads):
    """r_swNumThreads (SWRenderer.cpp:41, SWUtils.h:49-56) changes the per-thread float recurrences of RenderFinal and
    ApplyFog: osb200_opts.sw_threads replays them for 1, 8 and 32 workers (4 is the default every other test uses)."""
    solid, color = title_map.solid.copy(), title_map.color.copy()
    ref = oracles.Ref("sse2") if oracles.ref_available("sse2") else None
    for (w, h, fov) in ((800, 600, 30.0), (256, 256, 68.0)):
        scenes = [workloads.title_scene(w, h, fov)] + list(workloads.random_scenes(2, solid, w, h, fov, seed=22))
        with osb.Context(0, osb.NUM_SSE2, sw_threads=threads) as ctx:
            ctx.map_upload(solid, color)
            fbs, dps = ctx.render(np.stack(scenes), w, h, FOG, 128.0)
        if ref is not None:
            m = ref.map_from_arrays(solid, color)
            r = oracles.RefRenderer(ref, w, h, -1, threads, 0)
            r.set_map(m)
            r.set_fog(FOG, 128.0)
        for i, sc in enumerate(scenes):
            pfb, pdp = port.render(solid, color, sc, w, h, FOG, 128.0, osb.NUM_SSE2, threads, 1)
            assert _report(f"threads {threads} {w}x{h} #{i} vs port", fbs[i], dps[i], pfb, pdp) == (1.0, 1.0)
            if ref is not None:
                rfb, rdp = r.render(sc)
                assert _report(f"threads {threads} {w}x{h} #{i} vs ref", fbs[i], dps[i], rfb, rdp) == (1.0, 1.0)
        if ref is not None:
            r.set_map(None)
            r.close()
            ref.map_free(m)
    if ref is not None:
        ref.L.oref_renderer_set_threads(4)


def test_ceilings_hall(port, title_map):
    """a hall under a solid roof (undersides, pillars with several tops and bottoms per column): the ceilings list of
    BuildRle (SWMapRenderer.cpp:604-622), which outdoor cameras on Title.vxl never reach"""
    solid, color = title_map.solid.copy(), title_map.color.copy()
    rng = np.random.default_rng(3)
    x0, y0 = 200, 200
    for x in range(x0, x0 + 96):
        for y in range(y0, y0 + 96):
            m = (1 << 64) - 1
            m &= ~((1 << 5) - 1)
            m &= ~(((1 << 22) - 1) << 14)
            if rng.random() < 0.08:
                m |= ((1 << 6) - 1) << int(rng.integers(14, 30))
            if rng.random() < 0.05:
                m &= ~(((1 << 9) - 1) << 5)
            solid[x, y] = np.uint64(m)
            color[x, y, :] = (0x64000000 | rng.integers(0, 1 << 24, 64)).astype(np.uint32)
    w = h = 256
    scenes = []
    for k, (pitch, yaw) in enumerate([(0.0, 0.3), (-0.7, 2.0), (0.6, 4.0), (-1.2, 5.5), (1.3, 1.0)]):
        eye = np.array([x0 + 48.3, y0 + 47.6, 30.2 - k])
        d = np.array([np.cos(yaw) * np.cos(pitch), np.sin(yaw) * np.cos(pitch), np.sin(pitch)])
        scenes.append(workloads.look_at_scene(eye, eye + d, (0, 0, -1), 68.0, w, h))
    for mode in ("ieee", "sse2"):
        with osb.Context(0, NUMERICS[mode], sw_threads=4) as ctx:
            ctx.map_upload(solid, color)
            fbs, dps = ctx.render(np.stack(scenes), w, h, FOG, 128.0)
        for i, sc in enumerate(scenes):
            rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, NUMERICS[mode], 4, 1)
            assert _report(f"hall {mode} #{i}", fbs[i], dps[i], rfb, rdp) == (1.0, 1.0)
