"""CPU: pins the oracle restatement of the terrain pass + fog (oracle/orc_terrain.c) against the unmodified
reference SWRenderer (oracle/_ref) in its three numeric variants: V-SSE2 (default feature level
of the x86-64 build), V-SSE (same binary, SWFeatureLevel::SSE: rcpss + scalar shading) and V-IEEE (Draw/*.cpp
built with -U__SSE__ -U__SSE2__).  Frames and depth buffers must be bit-identical."""
import numpy as np
import pytest

import oracles
from openspades_b200 import workloads

pytestmark = pytest.mark.skipif(not oracles.ref_available(), reason="oracle/_ref not built and no /root/reference")

FOG = (0.1, 0.1, 0.1)
VARIANTS = {"sse2": ("sse2", -1, 2), "sse": ("sse2", 1, 1), "ieee": ("ieee", -1, 0)}  # lib, feature level, ORC_NUM_*


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_arrays(port):
    return port.vxl_load(open(oracles.TITLE_VXL, "rb").read())


def _compare(ref_lib, level, num, solid, color, scenes, w, h, threads, port, fog=FOG, under=1):
    ref = oracles.Ref(ref_lib)
    m = ref.map_from_arrays(solid, color)
    r = oracles.RefRenderer(ref, w, h, level, threads, under if under > 1 else 0)
    r.set_map(m)
    r.set_fog(fog, 128.0)
    depth = np.zeros((h, w), np.float32)  # the reference never clears its depth buffer: carry ours along too
    for i, sc in enumerate(scenes):
        wfb, wdp = r.render(sc)
        fb, dp = port.render(solid, color, sc, w, h, fog, 128.0, num, threads, under, depth=depth)
        assert np.array_equal(fb & 0xFFFFFF, wfb & 0xFFFFFF), (i, ((fb & 0xFFFFFF) == (wfb & 0xFFFFFF)).mean())
        assert np.array_equal(dp.view(np.uint32), wdp.view(np.uint32)), i
        if num == 2:
            assert np.array_equal(fb, wfb)  # including the junk X byte of the SSE2 fog path
    r.set_map(None)
    r.close()
    ref.map_free(m)
    ref.L.oref_renderer_set_undersampling(0)  # process-wide cvar: back to the default


@pytest.mark.parametrize("variant", ["sse2", "sse", "ieee"])
def test_title_scene_all_variants(port, title_arrays, variant):
    lib, level, num = VARIANTS[variant]
    solid, color = title_arrays
    _compare(lib, level, num, solid, color, [workloads.title_scene(800, 600, 30.0)], 800, 600, 4, port)


@pytest.mark.parametrize("variant", ["sse2", "ieee"])
def test_random_poses_256(port, title_arrays, variant):
    lib, level, num = VARIANTS[variant]
    solid, color = title_arrays
    scenes = list(workloads.random_scenes(12, solid, 256, 256, 68.0, seed=4321))
    _compare(lib, level, num, solid, color, scenes, 256, 256, 4, port, fog=(0.8, 1.0, 1.0))


def test_thread_count_changes_the_frame_and_the_port_follows(port, title_arrays):
    """r_swNumThreads is part of the result (per-thread float recurrences): 4 vs 8 threads differ, port tracks both."""
    solid, color = title_arrays
    sc = workloads.title_scene(1920, 1080, 68.0)
    a, _ = port.render(solid, color, sc, 1920, 1080, FOG, 128.0, 2, 4, 1)
    b, _ = port.render(solid, color, sc, 1920, 1080, FOG, 128.0, 2, 8, 1)
    assert not np.array_equal(a, b)
    _compare("sse2", -1, 2, solid, color, [sc], 1920, 1080, 8, port)


def test_synthetic_map_and_eye_inside_solid(port):
    solid, color = workloads.synthetic_map(7)
    scenes = list(workloads.random_scenes(4, solid, 256, 256, 68.0, seed=9))
    buried = workloads.title_scene(256, 256, 68.0)
    buried["origin"] = (100.5, 100.5, 62.5)
    scenes.insert(2, buried)  # keeps the previous frame's depth, clear colour fogged over it
    _compare("sse2", -1, 2, solid, color, scenes, 256, 256, 4, port)


@pytest.mark.parametrize("under", [2, 4])
@pytest.mark.parametrize("variant", ["sse2", "ieee"])
def test_undersampling_port_follows_the_cvar(port, title_arrays, under, variant):
    """r_swUndersampling 2 / 4 (SWMapRenderer.cpp:37,1110-1112,889-911,987-1009): fewer yaw lines, pixels replicated
    horizontally; 1, 4 and 8 worker threads (the replication interacts with the per-thread start columns)."""
    lib, level, num = VARIANTS[variant]
    solid, color = title_arrays
    for (w, h, fov) in ((800, 600, 30.0), (256, 256, 68.0), (320, 200, 68.0)):
        scenes = [workloads.title_scene(w, h, fov)] + list(workloads.random_scenes(2, solid, w, h, fov, seed=31))
        for threads in (4, 1, 8):
            _compare(lib, level, num, solid, color, scenes, w, h, threads, port, fog=(0.8, 1.0, 1.0), under=under)
