"""GPU parity of GameMap::CastRay (osb200_castray) and of the ambient-occlusion volume (osb200_ambient_*) through
the C ABI against the oracle restatement and, where oracle/_ref travelled, the reference's own GameMap /
GLAmbientShadowRenderer.  Bar: hit flags and blocks bit-exact (integer work); AO terms are written to be
bit-identical floats (the north_star's 1e-4 relative bar for the radiosity/ambient item is checked as well)."""
import numpy as np
import pytest

import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_map():
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    return m.solid.copy(), m.color.copy()


def _rays(n, seed):
    v0, d = workloads.random_rays(n, seed=seed)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    v0[: n // 8, 2] -= 40.0           # origins above the map
    v0[n // 8: n // 4, 2] += 30.0     # and below the water floor
    ev0, ed = workloads.edge_rays()
    keep = np.isfinite(ed).all(axis=1)  # GameMap::CastRay asserts its arguments are not NaN (GameMap.cpp:212-218) and an
    ev0, ed = ev0[keep], ed[keep]       # infinite direction makes v1 NaN / infinite: outside the function's contract
    return np.concatenate([v0, ev0]).astype(np.float32), np.concatenate([d.astype(np.float32), ed])


def _ao_rays(seed):
    """Same recipe as the reference's constructor: Normalize(U[0,1)^3) + 0.01"""
    rng = np.random.default_rng(seed)
    v = rng.random((16, 3), np.float32)
    return (v / np.linalg.norm(v, axis=1, keepdims=True) + np.float32(0.01)).astype(np.float32)


@pytest.mark.parametrize("length", [16.0, 128.0, 3.5, 0.0])
def test_castray_matches_oracle(port, title_map, length):
    solid, color = title_map
    v0, d = _rays(1 << 20, 77)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        got = ctx.castray(v0, d, length)
        assert ctx.castray(v0[:0], d[:0], length).shape == (0, 4)  # empty batch
    want = port.castray(solid, v0, d, length, threads=16)
    assert np.array_equal(got, want)


def test_castray_matches_reference_and_sees_edits(title_map):
    if not oracles.ref_available():
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("sse2")
    solid, color = title_map[0].copy(), title_map[1].copy()
    m = ref.map_from_arrays(solid, color)
    v0, d = _rays(1 << 18, 3)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        assert np.array_equal(ctx.castray(v0, d, 32.0), ref.castray(m, v0, d, 32.0, threads=16))
        rng = np.random.default_rng(8)
        for _ in range(2000):
            x, y, z = int(rng.integers(0, 512)), int(rng.integers(0, 512)), int(rng.integers(0, 63))
            was = (int(solid[x, y]) >> z) & 1
            ref.map_set(m, x, y, z, not was, 0x64112233)
            solid[x, y] = np.uint64(int(solid[x, y]) ^ (1 << z))
            color[x, y, z] = 0x64112233
            ctx.map_mark_dirty(x, y, z)
        assert ctx.map_flush(solid, color) > 0
        assert np.array_equal(ctx.castray(v0, d, 32.0), ref.castray(m, v0, d, 32.0, threads=16))
    ref.map_free(m)


def test_ao_evaluate_boxes_match_oracle(port, title_map):
    solid, color = title_map
    rays = _ao_rays(1)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        ctx.ambient_set_rays(rays)
        assert np.array_equal(ctx.ambient_rays(), rays)
        for lo, hi in (((250, 250, 20), (265, 262, 50)), ((504, 505, -2), (535, 518, 15)), ((-8, -6, 48), (7, 7, 65)),
                       ((100, 300, 0), (131, 301, 63)), ((7, 9, 11), (7, 9, 11))):
            got = ctx.ambient_evaluate_box(lo, hi)
            want = port.ao_evaluate_box(solid, rays, lo, hi, threads=16)
            assert np.abs(got - want).max() <= REL_TOL
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (lo, hi)


def test_ao_full_map_chunks_match_oracle(port, title_map):
    """BASELINE config 5's ambient half: every chunk dirty -> osb200_ambient_update(all); a spread of chunks
    (map corners = x/y wrap, sky and water-floor layers, terrain) is compared word for word."""
    solid, color = title_map
    rays = _ao_rays(2)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        ctx.ambient_set_rays(rays)
        assert ctx.ambient_chunk_info(3, 4, 1).tolist() == [1, 0, 0, 0, 15, 15, 15]
        assert ctx.ambient_update(-1) == 4096
        assert ctx.ambient_update(-1) == 0
        st = port.ao_state_init()
        chunks = [(0, 0, 0), (31, 31, 3), (0, 31, 2), (31, 0, 1), (16, 16, 2), (16, 16, 3), (10, 20, 1), (5, 27, 0),
                  (20, 9, 2), (15, 15, 3)]
        for c in chunks:
            want = port.ao_update_chunk(solid, rays, st, *c)
            got = ctx.ambient_read_chunk(*c)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), c
            assert ctx.ambient_chunk_info(*c)[0] == 0
        # size-independent properties over the whole volume: solid voxels are (0, 0), air flags are 1, AO in [0, 1]
        full = np.stack([ctx.ambient_read_chunk(cx, 12, cz) for cx in range(0, 32, 3) for cz in range(4)])
        assert full[..., 0].min() >= 0 and full[..., 0].max() <= 1
        assert set(np.unique(full[..., 1]).tolist()) <= {0.0, 1.0}
        assert np.all(full[..., 0][full[..., 1] == 0] == 0)


def test_ao_matches_reference_instance(title_map):
    """The unmodified GLAmbientShadowRenderer with ITS random rays handed to the GPU path."""
    if not oracles.ref_available():
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("sse2")
    solid, color = title_map
    m = ref.map_from_arrays(solid, color)
    ao = oracles.RefAmbient(ref, m)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        ctx.ambient_set_rays(ao.rays())
        ctx.ambient_update(-1)
        for c in ((16, 16, 2), (0, 31, 0), (31, 0, 3), (8, 8, 1)):
            ao.update_chunk(*c)
            assert np.array_equal(ctx.ambient_read_chunk(*c).view(np.uint32), ao.read_chunk(*c).view(np.uint32)), c
    ao.close()
    ref.map_free(m)


def test_ao_incremental_edits_match_oracle(port):
    """Frames of random dig/build edits -> mark_dirty + flush -> ambient update (budgeted and drain frames): the dirty
    boxes and the partially re-evaluated chunks equal a host mirror driven by the oracle."""
    solid, color = workloads.synthetic_map(3)
    rays = _ao_rays(4)
    rng = np.random.default_rng(99)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        ctx.ambient_set_rays(rays)
        assert ctx.ambient_update(-1) == 4096
        state = np.zeros((4096, 7), np.int32)
        solid0 = solid.copy()
        mirror = {}
        for frame in range(6):
            for _ in range(40):
                x, y = int(rng.integers(200, 330)), int(rng.integers(200, 330))
                col = int(solid[x, y])
                top = (col & -col).bit_length() - 1
                if rng.random() < 0.5 and top < 62:
                    z = top
                    solid[x, y] &= ~(np.uint64(1) << np.uint64(z))
                else:
                    z = max(top - 1, 0)
                    solid[x, y] |= np.uint64(1) << np.uint64(z)
                    color[x, y, z] = 0x64000000 | int(rng.integers(0, 1 << 24))
                ctx.map_mark_dirty(x, y, z)
                port.ao_invalidate(state, x, y, z)
            with pytest.raises(osb.OsbError, match="map_flush"):
                ctx.ambient_update(-1)
            assert ctx.map_flush(solid, color) > 0
            dirty = np.nonzero(state[:, 0])[0]
            assert len(dirty) > 0
            for i in dirty[:: max(1, len(dirty) // 16)]:
                cx, cy, cz = (i // 4) % 32, (i // 4) // 32, i % 4
                assert ctx.ambient_chunk_info(cx, cy, cz).tolist() == state[i].tolist()
            eye = (float(rng.integers(200, 330)), float(rng.integers(200, 330)), 20.0)
            n = ctx.ambient_update(8 if frame % 2 == 0 else -1, eye)
            if frame % 2 == 0:
                ex, ey, ez = int(eye[0]) >> 4, int(eye[1]) >> 4, int(eye[2]) >> 4
                near = [i for i in dirty if not (6 <= ((((i // 4) % 32) - ex) & 31) <= 26
                                                 or 6 <= ((((i // 4) // 32) - ey) & 31) <= 26
                                                 or (i % 4) - ez >= 6 or (i % 4) - ez <= -6)]
                taken = (near if near else list(dirty))[:8]
                assert n == len(taken)
            else:
                assert n == len(dirty)
                taken = list(dirty)
            for i in taken:  # every taken chunk is mirrored (a chunk keeps older values outside its dirty box)
                i = int(i)
                c = ((i // 4) % 32, (i // 4) // 32, i % 4)
                if i not in mirror:  # what the chunk held after the initial full update
                    s0 = port.ao_state_init()
                    mirror[i] = port.ao_update_chunk(solid0, rays, s0, *c)
                port.ao_update_chunk(solid, rays, state, c[0], c[1], c[2], mirror[i])
                assert state[i, 0] == 0
            for i in taken[:: max(1, len(taken) // 24)]:
                i = int(i)
                c = ((i // 4) % 32, (i // 4) // 32, i % 4)
                assert np.array_equal(ctx.ambient_read_chunk(*c).view(np.uint32), mirror[i].view(np.uint32)), (frame, c)
        assert ctx.ambient_update(-1) == 0
