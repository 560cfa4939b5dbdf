"""GPU parity of the VXL stream path (osb200_map_load_vxl / osb200_map_save_vxl through the C ABI) against the C
restatement of GameMap::Load / GameMap::Save and, when present, the unmodified reference.  Bar: bit-exact."""
import numpy as np
import pytest

import oracles
import vxl_cases
import openspades_b200 as osb
from openspades_b200 import workloads

pytestmark = pytest.mark.gpu
FOG = (0.8, 1.0, 1.0)


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


def test_device_load_of_title_matches_port_and_reference(port):
    data = open(oracles.TITLE_VXL, "rb").read()
    ws, wc = port.vxl_load(data)
    with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:
        gs, gc = ctx.map_load_vxl(data)
        assert np.array_equal(gs, ws) and np.array_equal(gc, wc)
        assert np.array_equal(ctx.map_read_solid(), ws)  # the resident (tiled) copy, read back in [x][y] order
        # the frame drawn from the device-decoded map is the frame drawn from the uploaded arrays
        sc = workloads.title_scene(256, 256, 68.0)
        fb1, dp1 = ctx.render(sc, 256, 256, FOG, 128.0)
        ctx.map_upload(ws, wc)
        fb2, dp2 = ctx.render(sc, 256, 256, FOG, 128.0)
        assert np.array_equal(fb1, fb2) and np.array_equal(dp1.view(np.uint32), dp2.view(np.uint32))
        assert ctx.map_save_vxl() == port.vxl_save(ws, wc)
    if oracles.ref_available("ieee"):
        ref = oracles.Ref("ieee")
        m = ref.map_load_vxl(data)
        rs, rc = ref.map_arrays(m)
        written = gc != 0
        assert np.array_equal(gs, rs) and np.array_equal(gc[written], rc[written])
        ref.map_free(m)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_device_save_and_load_of_synthetic_maps(port, seed):
    solid, color = vxl_cases.synthetic_arrays(seed, caves=seed != 3)
    want = port.vxl_save(solid, color)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        got = ctx.map_save_vxl()
        assert got == want
        gs, gc = ctx.map_load_vxl(got)
        ws, wc = port.vxl_load(want)
        assert np.array_equal(gs, ws) and np.array_equal(gc, wc)
        assert np.array_equal(gs, solid)          # geometry survives save -> load
        assert ctx.map_save_vxl() == got          # and the stream is a fixed point
    if oracles.ref_available("ieee"):
        ref = oracles.Ref("ieee")
        m = ref.map_from_arrays(solid, color)
        assert ref.map_save_vxl(m) == got
        ref.map_free(m)


def test_edits_after_a_device_load(port):
    """The arrays handed back by the device load are the host truth for GameMap::Set + osb200_map_flush."""
    data = open(oracles.TITLE_VXL, "rb").read()
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        solid, color = ctx.map_load_vxl(data)
        rng = np.random.default_rng(3)
        for _ in range(200):
            x, y = (int(v) for v in rng.integers(0, 512, 2))
            z = int(rng.integers(1, 62))
            solid[x, y] |= np.uint64(1) << np.uint64(z)
            color[x, y, z] = np.uint32(int(rng.integers(0, 1 << 24)) | (100 << 24))
            ctx.map_mark_dirty(x, y, z)
        with pytest.raises(osb.OsbError):
            ctx.map_save_vxl()  # unflushed edits
        ctx.map_flush(solid, color)
        assert ctx.map_save_vxl() == port.vxl_save(solid, color)


def test_truncated_and_malformed_streams_are_rejected():
    data = open(oracles.TITLE_VXL, "rb").read()
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        for bad in (data[:len(data) // 2], data[:-4], b"\x00\x00", b"\x00\x05\x03\x00" * 10):
            with pytest.raises(osb.OsbError):
                ctx.map_load_vxl(bad)
        sc = workloads.title_scene(64, 64, 68.0)
        # no map is resident after a failed load: the terrain pass is skipped like SetGameMap(null)
        fb, _ = ctx.render(sc, 64, 64, FOG, 128.0)
        ctx.map_load_vxl(data)  # and a good stream still loads afterwards
        fb2, _ = ctx.render(sc, 64, 64, FOG, 128.0)
        assert not np.array_equal(fb, fb2)


def test_short_stream_as_first_load_on_a_fresh_context():
    """a stream shorter than the column count leaves most column starts unwritten by the doubling rounds; they
    must read as 'unknown' (not as uninitialised device memory) and the load must fail cleanly, also as the very first
    load of a context, and leave the context usable."""
    data = open(oracles.TITLE_VXL, "rb").read()
    for bad in (b"\x00\x05\x03\x00" * 10, b"\x00\x3e\x3e\x00\x01\x02\x03\x04" * 1000, data[:4096]):
        with osb.Context(0, osb.NUM_IEEE) as ctx:
            with pytest.raises(osb.OsbError, match="malformed|truncated"):
                ctx.map_load_vxl(bad)
            ctx.map_load_vxl(data)
            ctx.sync()
