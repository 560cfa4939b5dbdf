"""The drop-in renderer through the reference's own interface.

openspades_b200/host/ref_adapter/B200Renderer.{h,cpp} is `class B200Renderer : public client::IRenderer, public
client::IGameMapListener`, constructed from a Handle<SWPort>, compiled against /root/reference/Sources by `make -C oracle
adapter` and linked with the reference translation units into oracle/_ref/libadapter_sse2.so (prebuilt here, it travels
to the GPU box).  oracle/ref_harness/adapter_test.cpp drives it and the unmodified draw::SWRenderer through the same
client::IRenderer* calls on one real GameMap (GameMap::Load of Title.vxl, GameMap::Set edits -> the real listener path in
both renderers, StartScene / RenderModel / AddLight / EndScene / FrameDone / ReadBitmap / Flip).  Here: what the two
ReadBitmap calls return must be identical."""
import ctypes
import os

import numpy as np
import pytest

import oracles
from openspades_b200 import workloads

pytestmark = pytest.mark.gpu

LIB = os.path.join(oracles.ORACLE_DIR, "_ref", "libadapter_sse2.so")


@pytest.fixture(scope="module")
def adapter():
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libadapter_sse2.so not built (needs the reference checkout: make -C oracle adapter)")
    L = ctypes.CDLL(LIB)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L.adapter_run.argtypes = [vp, ctypes.c_size_t, ci, ci, ci, ci, vp, ci, ci, ci, ctypes.c_uint32, vp, vp, vp, vp]
    L.adapter_last_error.restype = ctypes.c_char_p
    return L


def _run(L, w, h, scenes, edits, with_models, numerics=2, threads=4, seed=1):
    vxl = np.frombuffer(open(oracles.TITLE_VXL, "rb").read(), np.uint8)
    sc = np.ascontiguousarray(scenes)
    n = len(sc)
    a = np.zeros((n, h, w), np.uint32)
    b = np.zeros((n, h, w), np.uint32)
    info = np.zeros(4, np.int32)
    rc = L.adapter_run(vxl.ctypes.data, len(vxl), w, h, numerics, threads, sc.ctypes.data, n, edits, int(with_models), seed,
                       a.ctypes.data, b.ctypes.data, info.ctypes.data, None)
    assert rc == 0, L.adapter_last_error().decode()
    return a, b, info


@pytest.mark.parametrize("with_models", [False, True])
def test_same_frames_as_swrenderer_through_irenderer(adapter, with_models):
    port = oracles.Port()
    solid, _ = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    w, h = 320, 200
    scenes = np.concatenate([np.atleast_1d(workloads.title_scene(w, h, 68.0)),
                             workloads.random_scenes(5, solid, w, h, 68.0, seed=17, max_pitch_deg=35.0)])
    a, b, info = _run(adapter, w, h, scenes, edits=40, with_models=with_models)
    for i in range(len(scenes)):
        same = float((a[i] == b[i]).mean())
        assert same == 1.0, (i, same)
    assert np.all((a >> 24) == 0xFF)           # ReadBitmap forces alpha (SWRenderer.cpp:1093)
    assert info[2] == 40 * (len(scenes) - 1)   # the edits went through GameMap::Set
    assert 0 < info[0] <= 40 * 5               # ... and reached the renderer as dirty columns (edited + 4 neighbours)
    assert info[1] == len(scenes)              # one Swap per Flip; the pixel pointer changed every frame
    assert info[3] == 2                        # EndScene outside a scene raises in both renderers


def test_numerics_of_this_build(adapter):
    """the selection a maintainer would ship: B200Renderer::NumericsOfThisBuild() against SWRenderer at the feature level
    its own constructor detects on this host"""
    w, h = 256, 256
    scenes = np.stack([workloads.title_scene(w, h, 68.0)] * 2)
    a, b, _ = _run(adapter, w, h, scenes, edits=20, with_models=True, numerics=100)
    assert np.array_equal(a, b)


def test_edits_change_the_frame(adapter):
    """the same pose before and after digging next to the camera: the second frame differs from the first and still
    equals SWRenderer's (i.e. the listener path really re-uploaded the columns)"""
    w, h = 256, 256
    sc = workloads.title_scene(w, h, 68.0)
    scenes = np.stack([sc, sc, sc])
    a, b, info = _run(adapter, w, h, scenes, edits=300, with_models=False, seed=5)
    assert np.array_equal(a, b)
    assert not np.array_equal(a[0], a[1]) and not np.array_equal(a[1], a[2])
