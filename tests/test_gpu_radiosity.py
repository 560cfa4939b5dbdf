"""GPU parity of the shadow bitmap + radiosity (osb200_shadow_bitmap / osb200_radiosity_* through the C ABI)
against the oracle restatement and, where oracle/_ref travelled, the reference's own GLMapShadowRenderer /
GLRadiosityRenderer.  Bar (north_star): radiosity within 1e-4 relative; the kernels are written to be
bit-identical, so the tests assert equality of every float and every encoded word, and check the stated
tolerance separately."""
import numpy as np
import pytest

import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4  # BASELINE.json north_star
ABS_FLOOR = 1e-6


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_map():
    m = VoxelMap.load_file(oracles.TITLE_VXL)
    return m.solid.copy(), m.color.copy()


def _chunk_lo(i):
    return ((i // 4) % 32) * 16, ((i // 4) // 32) * 16, (i % 4) * 16


def test_shadow_bitmap_matches_oracle(port, title_map):
    solid, color = title_map
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        got = ctx.shadow_bitmap()
    assert np.array_equal(got, port.shadow_bitmap(solid, color))
    solid2, color2 = workloads.synthetic_map(5)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid2, color2)
        assert np.array_equal(ctx.shadow_bitmap(), port.shadow_bitmap(solid2, color2))


@pytest.mark.parametrize("mode", [1, 2])
def test_evaluate_boxes_match_oracle(port, title_map, mode):
    solid, color = title_map
    bm = port.shadow_bitmap(solid, color)
    with osb.Context(0, osb.NUM_IEEE, radiosity_mode=mode) as ctx:
        ctx.map_upload(solid, color)
        for lo, hi in (((250, 250, 20), (265, 262, 50)), ((504, 505, 0), (535, 518, 15)), ((-8, -6, 40), (7, 7, 63)),
                       ((100, 300, 0), (131, 301, 63)), ((7, 9, 11), (7, 9, 11))):
            got = ctx.radiosity_evaluate_box(lo, hi)
            want, _ = port.radiosity_box(bm, lo, hi, mode, want_enc=False)
            err = np.abs(got - want) / np.maximum(np.abs(want), ABS_FLOOR)
            assert err.max() <= REL_TOL
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (lo, hi)


@pytest.mark.parametrize("mode", [1, 2])
def test_full_map_update_matches_oracle(port, title_map, mode):
    """All 4096 chunks dirty (initial state) -> one osb200_radiosity_update(all); encoded words of a spread of
    chunks against the oracle, and full-volume sanity."""
    solid, color = title_map
    bm = port.shadow_bitmap(solid, color)
    with osb.Context(0, osb.NUM_IEEE, radiosity_mode=mode) as ctx:
        ctx.map_upload(solid, color)
        assert ctx.radiosity_chunk_info(3, 4, 1).tolist() == [1, 0, 0, 0, 15, 15, 15]
        assert np.all(ctx.radiosity_read_chunk(3, 4, 1) == 0x20080200)
        assert ctx.radiosity_update(-1) == 4096
        assert ctx.radiosity_update(-1) == 0
        assert ctx.radiosity_chunk_info(3, 4, 1)[0] == 0
        rng = np.random.default_rng(5)
        ids = list(rng.integers(0, 4096, 24)) + [0, 3, 4095, 4 * 31, 4 * 32 * 31 + 2]
        for i in ids:
            x0, y0, z0 = _chunk_lo(int(i))
            _, enc = port.radiosity_box(bm, (x0, y0, z0), (x0 + 15, y0 + 15, z0 + 15), mode, want_f32=False)
            got = ctx.radiosity_read_chunk(x0 // 16, y0 // 16, z0 // 16)
            assert np.array_equal(got, enc), i
            assert np.all(got >> 30 == 3)


def test_against_unmodified_reference(title_map):
    if not oracles.ref_available("sse2"):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("sse2")
    solid, color = title_map
    m = ref.map_from_arrays(solid, color)
    rr = oracles.RefRadiosity(ref, m, 1)
    with osb.Context(0, osb.NUM_IEEE, radiosity_mode=1) as ctx:
        ctx.map_upload(solid, color)
        assert np.array_equal(ctx.shadow_bitmap(), rr.shadow_bitmap())
        ctx.radiosity_update(-1)
        for (cx, cy, cz) in ((16, 16, 2), (0, 31, 0), (31, 0, 3), (12, 20, 1)):
            rr.update_chunk(cx, cy, cz)
            assert np.array_equal(ctx.radiosity_read_chunk(cx, cy, cz), rr.read_chunk(cx, cy, cz))
        lo, hi = (240, 250, 30), (260, 258, 50)
        assert np.array_equal(ctx.radiosity_evaluate_box(lo, hi).view(np.uint32), rr.evaluate_box(lo, hi).view(np.uint32))
    rr.close()
    ref.map_free(m)


def test_incremental_edits_match_oracle(port):
    """BASELINE config 5 in small: frames of random dig/build edits -> mark_dirty + flush -> radiosity update.
    After every frame the shadow bitmap, the set of dirty chunks with their boxes, and the words of every
    chunk taken so far must equal a host mirror driven by the oracle."""
    solid, color = workloads.synthetic_map(3)
    bm = port.shadow_bitmap(solid, color)
    state = np.zeros((4096, 7), np.int32)
    words = np.zeros(512 * 16, np.uint32)
    rng = np.random.default_rng(99)
    with osb.Context(0, osb.NUM_IEEE, radiosity_mode=1) as ctx:
        ctx.map_upload(solid, color)
        assert ctx.radiosity_update(-1) == 4096
        bm0 = bm.copy()
        mirror = {}
        for frame in range(8):
            for _ in range(100):
                x, y = int(rng.integers(0, 512)), int(rng.integers(0, 512))
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
                port.shadow_mark(words, x, y, z)
            with pytest.raises(osb.OsbError, match="map_flush"):
                ctx.shadow_bitmap()
            assert ctx.map_flush(solid, color) > 0
            port.shadow_update(solid, color, bm, words, state)
            assert np.array_equal(ctx.shadow_bitmap(), bm)
            dirty = np.nonzero(state[:, 0])[0]
            assert len(dirty) > 0
            for i in dirty[:: max(1, len(dirty) // 16)]:
                cx, cy, cz = (i // 4) % 32, (i // 4) // 32, i % 4
                assert ctx.radiosity_chunk_info(cx, cy, cz).tolist() == state[i].tolist()
            eye = (float(rng.integers(0, 512)), float(rng.integers(0, 512)), 20.0)
            n = ctx.radiosity_update(8 if frame % 2 == 0 else -1, eye)  # budgeted frames and drain frames
            if frame % 2 == 0:
                # which chunks were taken is our policy (near the eye first, index order): mirror it
                ex, ey, ez = int(eye[0]) >> 4, int(eye[1]) >> 4, int(eye[2]) >> 4
                near = [i for i in dirty if not (6 <= ((((i // 4) % 32) - ex) & 31) <= 26
                                                 or 6 <= ((((i // 4) // 32) - ey) & 31) <= 26
                                                 or (i % 4) - ez >= 6 or (i % 4) - ez <= -6)]
                taken = (near if near else list(dirty))[:8]
                assert n == len(taken)
            else:
                assert n == len(dirty)
                taken = list(dirty)
            for i in taken:
                i = int(i)
                x0, y0, z0 = _chunk_lo(i)
                if i not in mirror:  # words this chunk got from the initial full update
                    mirror[i] = port.radiosity_box(bm0, (x0, y0, z0), (x0 + 15, y0 + 15, z0 + 15), 1, want_f32=False)[1]
                b = state[i]
                lo = (x0 + b[1], y0 + b[2], z0 + b[3])
                hi = (x0 + b[4], y0 + b[5], z0 + b[6])
                # UpdateChunk recomputes the dirty box only; voxels outside it keep their older words (the
                # reference's +-6 invalidation is deliberately approximate along the sun diagonal)
                mirror[i][b[3]:b[6] + 1, b[2]:b[5] + 1, b[1]:b[4] + 1] = port.radiosity_box(bm, lo, hi, 1, want_f32=False)[1]
                state[i, 0] = 0
            for i in taken[:: max(1, len(taken) // 12)]:
                x0, y0, z0 = _chunk_lo(int(i))
                assert np.array_equal(ctx.radiosity_read_chunk(x0 // 16, y0 // 16, z0 // 16), mirror[int(i)]), (frame, i)
        assert ctx.radiosity_update(-1) == 0
        for i in sorted(mirror)[:: max(1, len(mirror) // 40)]:
            x0, y0, z0 = _chunk_lo(i)
            assert np.array_equal(ctx.radiosity_read_chunk(x0 // 16, y0 // 16, z0 // 16), mirror[i]), i
