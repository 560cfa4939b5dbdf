"""Pins the VXL stream restatements (oracle/orc_map.c: orc_vxl_load, orc_vxl_save) against GameMap::Load / GameMap::Save of
the unmodified reference (oracle/_ref), on the bundled map and on synthetic maps with caves and floating slabs."""
import numpy as np
import pytest

import oracles
import vxl_cases


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def ref():
    if not oracles.ref_available("ieee"):
        pytest.skip("oracle/_ref not present")
    return oracles.Ref("ieee")


def test_save_of_title_matches_reference_and_reloads(port, ref):
    data = open(oracles.TITLE_VXL, "rb").read()
    m = ref.map_load_vxl(data)
    want = ref.map_save_vxl(m)
    solid, color = port.vxl_load(data)
    got = port.vxl_save(solid, color)
    assert got == want
    # Save keeps the geometry and every visible colour: loading the saved stream gives the same solid map and a
    # stream that saves to itself
    s2, c2 = port.vxl_load(got)
    assert np.array_equal(s2, solid)
    assert port.vxl_save(s2, c2) == got
    ref.map_free(m)


def _spans(stream: bytes):
    """number of span headers in a stream (walks it like GameMap::Load)"""
    b = np.frombuffer(stream, np.uint8)
    pos = n = 0
    while pos < len(b):
        n += 1
        pos += 4 * int(b[pos]) if b[pos] else 4 * (int(b[pos + 2]) - int(b[pos + 1]) + 2)
    return n


@pytest.mark.parametrize("seed", [1, 2])
def test_save_and_load_of_synthetic_maps_match_reference(port, ref, seed):
    solid, color = vxl_cases.synthetic_arrays(seed)
    m = ref.map_from_arrays(solid, color)
    want = ref.map_save_vxl(m)
    assert port.vxl_save(solid, color) == want
    m2 = ref.map_load_vxl(want)
    ws, wc = ref.map_arrays(m2)
    gs, gc = port.vxl_load(want)
    written = gc != 0  # voxels the stream never names keep GameMap's random default colours (GameMap.cpp:36-48)
    assert np.array_equal(gs, ws) and np.array_equal(gc[written], wc[written])
    assert written.sum() >= (len(want) - 4 * _spans(want)) // 4 * 0.99  # about one written voxel per colour word
    assert np.array_equal(gs, solid)  # geometry survives the round trip
    ref.map_free(m)
    ref.map_free(m2)
