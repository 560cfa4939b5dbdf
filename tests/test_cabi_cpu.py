"""CPU: the C-ABI library loads, exports every symbol include/osb200.h declares, and refuses to compute
without a CUDA device (there is no CPU fallback in the product).  No kernel is launched here."""
import ctypes
import os
import re

import numpy as np
import pytest

import openspades_b200 as osb
from openspades_b200 import binding, hostlib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "osb200.h")).read()
    return sorted(set(re.findall(r"OSB200_API[^;(]*?\b(osb200_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = binding.load_library()
    names = _declared_symbols()
    assert len(names) >= 25 and "osb200_render" in names and "osb200_radiosity_update" in names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_struct_layouts_match_the_reference_types():
    # GameMap::RayCastResult (GameMap.h:150-156) is 40 bytes; osb200_scene is 64
    assert binding.RAYHIT_DTYPE.itemsize == 40 and binding.SCENE_DTYPE.itemsize == 64
    assert binding.RAYHIT_DTYPE.fields["hitPos"][1] == 4 and binding.RAYHIT_DTYPE.fields["hitBlock"][1] == 16
    assert binding.RAYHIT_DTYPE.fields["normal"][1] == 28


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="a CUDA device is present")
def test_no_gpu_fails_loudly():
    with pytest.raises(osb.OsbError, match="no CUDA device|no CPU fallback"):
        osb.Context(device=0)
    with pytest.raises(osb.OsbError, match="no CUDA device|no CPU fallback"):
        hostlib.B200Renderer(64, 64)


def test_sse_calibration_is_host_only_and_consistent():
    rcp, rb, rsq, sb = osb.calibrate_sse()
    assert 8 <= rb <= 16 and 8 <= sb <= 16 and len(rcp) == 1 << rb and len(rsq) == 2 << sb
    # table entry 0 = rcpss(1.0): within 1.5 * 2^-12 of 1 (Intel SDM bound for RCPSS)
    one = rcp[:1].view(np.float32)[0]
    assert abs(one - 1.0) <= 1.5 * 2.0 ** -12


def test_host_voxelmap_set_and_listener_contract():
    m = hostlib.VoxelMap()
    assert m.solid.shape == (512, 512) and m.color.shape == (512, 512, 64)
    before = int(m.solid[5, 6])
    m.set(5, 6, 10, True, 0x64112233)
    assert int(m.solid[5, 6]) == before | (1 << 10) and int(m.color[5, 6, 10]) == 0x64112233
    m.set(5, 6, 10, False, 0)
    assert int(m.solid[5, 6]) == before
    with pytest.raises(osb.OsbError):
        m.set(512, 0, 0, True, 0)


def test_host_vxl_loader_rejects_malformed_streams():
    """VoxelMap::Load (GameMap::Load, GameMap.cpp:519-596) takes untrusted network input: every malformed stream must end in
    an exception, never in a write outside the 64-entry column or an endless walk (z / chunk checks)."""
    import oracles
    from openspades_b200.binding import OsbError
    data = open(oracles.TITLE_VXL, "rb").read()
    VoxelMap = hostlib.VoxelMap
    good = VoxelMap.load(data)
    ps, pc = oracles.Port().vxl_load(data)
    assert np.array_equal(good.solid, ps)
    named = ps[:, :, None] >> np.arange(64, dtype=np.uint64) & np.uint64(1) != 0
    surface_equal = (good.color == pc) | ~named
    assert surface_equal.mean() > 0.99  # colours the stream names agree (the rest is the reference's random default)
    col = lambda *spans: b"".join(bytes(s) for s in spans)  # noqa: E731
    bad_streams = {
        "truncated half": data[:len(data) // 2],
        "two bytes": b"\x00\x00",
        "z beyond 63": col([0, 60, 70, 0] + [1, 2, 3, 4] * 11) * 10,
        "negative topStart": col([0, 0x90, 3, 0] + [0] * 4 * 120) * 4,
        "negative chunks": col([0x80, 0, 0, 0, 1, 2, 3, 4]) * 8,
        "colour run crosses header": col([2, 0, 5, 0, 9, 9, 9, 9, 0, 20, 20, 20, 1, 1, 1, 1]) * 8,
        "endless spans": col([1, 0, -1 & 0xff, 0]) * 4096,
    }
    for name, bad in bad_streams.items():
        with pytest.raises(OsbError):
            VoxelMap.load(bad)


def test_reference_side_adapter_links_and_fails_loudly_without_a_gpu():
    """oracle/_ref/libadapter_sse2.so = openspades_b200/host/ref_adapter/B200Renderer.cpp (class B200Renderer : public
    client::IRenderer, public client::IGameMapListener) compiled against the reference headers and linked with the reference
    translation units.  Without a CUDA device constructing it must raise the reference's own exception type with the
    library's "no CPU fallback" message (the frames themselves are compared on the GPU: tests/test_gpu_adapter.py)."""
    import oracles
    from openspades_b200 import workloads
    lib = os.path.join(oracles.ORACLE_DIR, "_ref", "libadapter_sse2.so")
    if not os.path.exists(lib):
        pytest.skip("adapter not built (needs the reference checkout: make -C oracle adapter)")
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present: covered by tests/test_gpu_adapter.py")
    except ImportError:
        pass
    L = ctypes.CDLL(lib)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L.adapter_run.argtypes = [vp, ctypes.c_size_t, ci, ci, ci, ci, vp, ci, ci, ci, ctypes.c_uint32, vp, vp, vp, vp]
    L.adapter_last_error.restype = ctypes.c_char_p
    vxl = np.frombuffer(open(oracles.TITLE_VXL, "rb").read(), np.uint8)
    sc = np.ascontiguousarray(np.atleast_1d(workloads.title_scene(64, 64, 68.0)))
    a = np.zeros((1, 64, 64), np.uint32)
    b = np.zeros((1, 64, 64), np.uint32)
    info = np.zeros(4, np.int32)
    rc = L.adapter_run(vxl.ctypes.data, len(vxl), 64, 64, 2, 4, sc.ctypes.data, 1, 0, 0, 1, a.ctypes.data, b.ctypes.data,
                       info.ctypes.data, None)
    assert rc == -1
    msg = L.adapter_last_error().decode()
    assert "no CPU fallback" in msg and "B200Renderer" in msg
