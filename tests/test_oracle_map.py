"""CPU: the C restatement (oracle/liboracle.so) and the product's host VoxelMap against the unmodified
reference (oracle/_ref) for VXL decode, BuildRle records and CastRay2."""
import numpy as np
import pytest

import oracles
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.skipif(not oracles.ref_available(), reason="oracle/_ref not built and no /root/reference")


@pytest.fixture(scope="module")
def ref():
    return oracles.Ref("sse2")


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title(ref):
    data = open(oracles.TITLE_VXL, "rb").read()
    m = ref.map_load_vxl(data)
    solid, color = ref.map_arrays(m)
    yield data, m, solid, color
    ref.map_free(m)


def test_vxl_decode_matches_reference(ref, port, title):
    data, m, solid, color = title
    ps, pc = port.vxl_load(data)
    assert np.array_equal(ps, solid)
    assert int(np.unpackbits(solid.view(np.uint8)).sum()) == 3778939  # Title.vxl
    written = pc != 0
    assert written.sum() > 250000
    assert np.array_equal(pc[written], color[written])
    # product host loader (host/VoxelMap.cpp)
    hm = VoxelMap.load(data)
    assert np.array_equal(hm.solid, solid)
    assert np.array_equal(hm.color, pc)
    with pytest.raises(Exception):
        VoxelMap.load(data[:1000])
    with pytest.raises(ValueError):
        port.vxl_load(data[:1000])


def test_rle_records_match_reference(ref, port, title):
    _, m, solid, _ = title
    r = oracles.RefRenderer(ref, 64, 64)
    r.set_map(m)
    rng = np.random.default_rng(5)
    cols = [(0, 0), (511, 511), (0, 511), (511, 0), (256, 256)] + [tuple(rng.integers(0, 512, 2)) for _ in range(2000)]
    for x, y in cols:
        a = r.rle(int(x), int(y))
        b = port.build_rle(solid, int(x), int(y))
        assert np.array_equal(a, b), (x, y)
    r.set_map(None)
    r.close()


@pytest.mark.parametrize("max_steps", [256, 64, 12, 1])
def test_castray2_port_matches_reference(ref, port, title, max_steps):
    _, m, solid, _ = title
    v0, d = workloads.random_rays(200000, seed=1234 + max_steps)
    a = ref.castray2(m, v0, d, max_steps, threads=8)
    b = port.castray2(solid, v0, d, max_steps, threads=8)
    assert a.tobytes() == b.tobytes()
    assert 0.05 < a["hit"].mean() < 1.0 and a["startSolid"].sum() > 1000


def test_castray2_edge_rays(ref, port, title):
    _, m, solid, _ = title
    v0, d = workloads.edge_rays()  # incl. zero, underflowing, infinite and NaN directions
    a = ref.castray2(m, v0, d, 300)
    b = port.castray2(solid, v0, d, 300)
    assert a.tobytes() == b.tobytes()


def test_castray2_synthetic_map(ref, port):
    solid, color = workloads.synthetic_map(7)
    m = ref.map_from_arrays(solid, color)
    v0, d = workloads.random_rays(100000, seed=99)
    a = ref.castray2(m, v0, d, 256, threads=8)
    b = port.castray2(solid, v0, d, 256, threads=8)
    ref.map_free(m)
    assert a.tobytes() == b.tobytes()
