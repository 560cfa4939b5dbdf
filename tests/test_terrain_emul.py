"""CPU lock-step emulation of the terrain kernels against the oracle (no GPU needed).

tests/native/terrain_emul.cpp compiles openspades_b200/csrc/terrain_core.cuh -- the per-lane code of k_line_setup,
k_walk_lines and k_gather_fog -- with g++ and runs it in the kernels' launch structure.  What runs on the B200 is the
same header under nvcc; the -m gpu tests repeat these comparisons through the C ABI.  Here the frames must already be
bit-identical to the oracle restatement (which tests/test_oracle_terrain.py pins to the unmodified reference)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import oracles
from openspades_b200 import workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NATIVE = os.path.join(ROOT, "tests", "native")
LIB = os.path.join(NATIVE, "_build", "libterrain_emul.so")
CSRC = os.path.join(ROOT, "openspades_b200", "csrc")
FOG = (0.8, 1.0, 1.0)


def _build():
    srcs = [os.path.join(NATIVE, "terrain_emul.cpp"), os.path.join(CSRC, "sse_calibrate.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("terrain_core.cuh", "hd.cuh", "numerics.cuh", "libm_f32.cuh")]
    if os.path.exists(LIB) and all(os.path.getmtime(d) <= os.path.getmtime(LIB) for d in deps):
        return
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-o", LIB] + srcs)


class Emul:
    def __init__(self):
        _build()
        self.L = ctypes.CDLL(LIB)
        vp, ci = ctypes.c_void_p, ctypes.c_int
        self.L.emul_render.argtypes = [vp, vp, vp, ci, ci, ci, vp, ctypes.c_float, ci, ci, ci, ci, ci, vp, vp, vp, ci]
        self.L.emul_set_coop.argtypes = [ci]
        self.key = 0

    def set_walker(self, mode):
        """0 = lane-per-line walker (k_walk_lines), 1 = cooperative walker (k_walk_coop), 2 = its sequential fallback"""
        self.L.emul_set_coop(mode)

    def render(self, solid, color, scenes, w, h, fog, fog_dist, numerics, threads=4, under=1, no_map=False, depth=None,
               new_map=True):
        sc = np.ascontiguousarray(np.atleast_1d(scenes))
        n = len(sc)
        fb = np.zeros((n, h, w), np.uint32)
        dp = np.zeros((n, h, w), np.float32) if depth is None else depth
        fc = np.asarray(fog, np.float32)
        st = np.zeros(3, np.uint64)
        if new_map:
            self.key += 1
        rc = self.L.emul_render(solid.ctypes.data, color.ctypes.data, sc.ctypes.data, n, w, h, fc.ctypes.data, fog_dist,
                                numerics, threads, under, int(no_map), 0, fb.ctypes.data, dp.ctypes.data, st.ctypes.data,
                                self.key)
        assert rc == 0, rc
        return fb, dp, st


@pytest.fixture(scope="module", params=[0, 1], ids=["lane_per_line", "cooperative"])
def emul(request):
    e = Emul()
    e.set_walker(request.param)
    yield e
    e.set_walker(0)


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_map(port):
    return port.vxl_load(open(oracles.TITLE_VXL, "rb").read())


def _same(tag, fb, dp, rfb, rdp):
    a = float(((fb & 0xFFFFFF) == (rfb & 0xFFFFFF)).mean())
    b = float((dp.view(np.uint32) == rdp.view(np.uint32)).mean())
    if a < 1 or b < 1:
        print(f"{tag}: rgb identical {a:.6f} depth identical {b:.6f}")
    return a, b


@pytest.mark.parametrize("numerics", [0, 1, 2])
@pytest.mark.parametrize("size", [(256, 256, 68.0), (800, 600, 30.0)])
def test_title_scene(emul, port, title_map, numerics, size):
    solid, color = title_map
    w, h, fov = size
    sc = workloads.title_scene(w, h, fov)
    fb, dp, st = emul.render(solid, color, sc, w, h, FOG, 128.0, numerics)
    rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, numerics, 4, 1)
    assert _same(f"title {w}x{h} num {numerics}", fb[0], dp[0], rfb, rdp) == (1.0, 1.0)
    if numerics == 2:  # full 32-bit word including the reference's junk X byte
        assert np.array_equal(fb[0], rfb)
    assert 0 < st[1] < st[0]  # some lines are pitch-culled, some survive


@pytest.mark.parametrize("numerics", [0, 2])
def test_random_poses_synthetic_map(emul, port, numerics):
    """seeded poses with pitch +-60 deg: covers the yawRange = pi and the pole-visible 2 pi case, ceilings (caves)."""
    solid, color = workloads.synthetic_map(7)
    w = h = 256
    scenes = workloads.random_scenes(12, solid, w, h, 68.0, seed=4321)
    fb, dp, _ = emul.render(solid, color, scenes, w, h, FOG, 128.0, numerics)
    for i in range(len(scenes)):
        rfb, rdp = port.render(solid, color, scenes[i], w, h, FOG, 128.0, numerics, 4, 1)
        assert _same(f"pose {i} num {numerics}", fb[i], dp[i], rfb, rdp) == (1.0, 1.0)


@pytest.mark.parametrize("under", [2, 4])
@pytest.mark.parametrize("size", [(256, 256, 68.0), (320, 200, 68.0)])
def test_undersampling(emul, port, title_map, under, size):
    """r_swUndersampling 2 / 4 (SWMapRenderer.cpp:37,1110-1112; store replication :889-911,987-1009)"""
    solid, color = title_map
    w, h, fov = size
    scenes = [workloads.title_scene(w, h, fov)] + list(workloads.random_scenes(2, solid, w, h, fov, seed=5))
    for i, sc in enumerate(scenes):
        fb, dp, _ = emul.render(solid, color, sc, w, h, FOG, 128.0, 2, under=under, new_map=(i == 0))
        rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, 2, 4, under)
        assert _same(f"under {under} {w}x{h} #{i}", fb[0], dp[0], rfb, rdp) == (1.0, 1.0)


@pytest.mark.parametrize("threads", [1, 3, 8, 32])
def test_sw_threads(emul, port, title_map, threads):
    """r_swNumThreads changes the per-thread float recurrences of RenderFinal / ApplyFog (SWUtils.h:49-56)"""
    solid, color = title_map
    w, h = 800, 600
    sc = workloads.title_scene(w, h, 30.0)
    fb, dp, _ = emul.render(solid, color, sc, w, h, FOG, 128.0, 2, threads=threads)
    rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, 2, threads, 1)
    assert _same(f"threads {threads}", fb[0], dp[0], rfb, rdp) == (1.0, 1.0)


def test_eye_inside_solid_keeps_depth(emul, port, title_map):
    solid, color = title_map
    sc = workloads.title_scene(64, 64, 68.0)
    sc["origin"] = (256.5, 256.5, 62.5)
    depth = np.full((1, 64, 64), 7.5, np.float32)
    fb, dp, _ = emul.render(solid, color, sc, 64, 64, FOG, 128.0, 2, depth=depth)
    rdp0 = np.full((64, 64), 7.5, np.float32)
    rfb, rdp = port.render(solid, color, sc, 64, 64, FOG, 128.0, 2, 4, 1, depth=rdp0)
    assert np.array_equal(fb[0] & 0xFFFFFF, rfb & 0xFFFFFF) and np.array_equal(dp[0], rdp)


@pytest.mark.parametrize("numerics", [0, 2])
def test_ceilings_above_the_eye(emul, port, title_map, numerics):
    """a hall carved under a solid roof: the eye sees undersides (PosZ faces, the ceilings list of BuildRle,
    SWMapRenderer.cpp:604-622) and pillars with several tops / bottoms per column"""
    solid, color = title_map[0].copy(), title_map[1].copy()
    rng = np.random.default_rng(3)
    x0, y0 = 200, 200
    for x in range(x0, x0 + 96):
        for y in range(y0, y0 + 96):
            m = (1 << 64) - 1
            m &= ~(((1 << 5) - 1) << 0)           # z 0..4 air (sky above the roof)
            m &= ~(((1 << 22) - 1) << 14)         # z 14..35 air: the hall; roof z 5..13, floor z 36..63
            if rng.random() < 0.08:               # pillars with gaps
                m |= ((1 << 6) - 1) << int(rng.integers(14, 30))
            if rng.random() < 0.05:               # holes in the roof
                m &= ~(((1 << 9) - 1) << 5)
            solid[x, y] = np.uint64(m)
            color[x, y, :] = (0x64000000 | rng.integers(0, 1 << 24, 64)).astype(np.uint32)
    w = h = 256
    scenes = []
    for k, (pitch, yaw) in enumerate([(0.0, 0.3), (-0.7, 2.0), (0.6, 4.0), (-1.2, 5.5), (1.3, 1.0)]):
        eye = np.array([x0 + 48.3, y0 + 47.6, 30.2 - k])
        d = np.array([np.cos(yaw) * np.cos(pitch), np.sin(yaw) * np.cos(pitch), np.sin(pitch)])
        scenes.append(workloads.look_at_scene(eye, eye + d, (0, 0, -1), 68.0, w, h))
    for i, sc in enumerate(scenes):
        fb, dp, _ = emul.render(solid, color, sc, w, h, FOG, 128.0, numerics, new_map=(i == 0))
        rfb, rdp = port.render(solid, color, sc, w, h, FOG, 128.0, numerics, 4, 1)
        assert _same(f"hall {i} num {numerics}", fb[0], dp[0], rfb, rdp) == (1.0, 1.0)
        dark = ((rfb & 0xFFFFFF) != 0).mean()
        assert dark > 0.5


def test_cooperative_fallback_path(port, title_map):
    """a line with more spans than the cooperative walker records falls back to the sequential one inside the same warp"""
    e = Emul()
    e.set_walker(2)
    solid, color = title_map
    sc = workloads.title_scene(256, 256, 68.0)
    fb, dp, _ = e.render(solid, color, sc, 256, 256, FOG, 128.0, 2)
    e.set_walker(0)
    rfb, rdp = port.render(solid, color, sc, 256, 256, FOG, 128.0, 2, 4, 1)
    assert _same("coop fallback", fb[0], dp[0], rfb, rdp) == (1.0, 1.0)
