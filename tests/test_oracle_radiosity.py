"""CPU: pins the oracle restatement of the shadow bitmap + radiosity (oracle/orc_radiosity.c) against the
unmodified reference classes GLMapShadowRenderer / GLRadiosityRenderer (oracle/_ref, built in place from
/root/reference by oracle/Makefile with a do-nothing IGLDevice).  Bit-exact on every float and word."""
import numpy as np
import pytest

import oracles
from openspades_b200 import workloads

pytestmark = pytest.mark.skipif(not oracles.ref_available(), reason="oracle/_ref not built and no /root/reference")


@pytest.fixture(scope="module")
def ref():
    return oracles.Ref("sse2")


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title(ref):
    m = ref.map_load_vxl(open(oracles.TITLE_VXL, "rb").read())
    solid, color = ref.map_arrays(m)
    yield m, solid, color
    ref.map_free(m)


def test_shadow_bitmap_matches_reference(ref, port, title):
    m, solid, color = title
    rr = oracles.RefRadiosity(ref, m, 1)
    want = rr.shadow_bitmap()
    got = port.shadow_bitmap(solid, color)
    assert np.array_equal(got, want)
    assert (want >> 24).max() == 64 and ((want >> 7) & 1).any()
    rr.close()


@pytest.mark.parametrize("mode", [1, 2])
def test_evaluate_and_encode_match_reference(ref, port, title, mode):
    m, solid, color = title
    rr = oracles.RefRadiosity(ref, m, mode)
    bm = rr.shadow_bitmap()
    # boxes: map interior, x/y wrap-around corner, sky and water-floor slabs
    for lo, hi in (((250, 250, 20), (265, 262, 50)), ((504, 505, 0), (519, 518, 15)), ((-8, -6, 48), (7, 7, 63)),
                   ((100, 300, 0), (131, 301, 63))):
        want = rr.evaluate_box(lo, hi)
        f32, enc = port.radiosity_box(bm, lo, hi, mode)
        assert np.array_equal(f32.view(np.uint32), want.view(np.uint32)), (lo, hi)
        assert np.isfinite(want).all() and want[..., :3].min() >= 0  # the reference's own SPAsserts
        flat = want.reshape(-1, 12)
        rng = np.random.default_rng(1)
        for i in rng.integers(0, len(flat), 200):
            for k in range(4):
                assert rr.encode(flat[i, k * 3:k * 3 + 3]) == int(enc.reshape(-1, 4)[i, k])
    # encode edge values
    for v in ((0, 0, 0), (1, 1, 1), (-1, -1, -1), (5, -5, 0.25), (1e-8, -1e-8, 0.999), (0.5, -0.25, 1e30)):
        assert rr.encode(v) == port.radiosity_encode(v, mode), v
    rr.close()


def test_update_chunk_words_match_reference(ref, port, title):
    m, solid, color = title
    rr = oracles.RefRadiosity(ref, m, 1)
    bm = rr.shadow_bitmap()
    assert rr.num_dirty() == 4096  # initial state: everything dirty (GLRadiosityRenderer.h:56-59)
    for (cx, cy, cz) in ((16, 16, 2), (0, 31, 0), (31, 0, 3)):
        rr.update_chunk(cx, cy, cz)
        want = rr.read_chunk(cx, cy, cz)
        lo = (cx * 16, cy * 16, cz * 16)
        _, enc = port.radiosity_box(bm, lo, (lo[0] + 15, lo[1] + 15, lo[2] + 15), 1, want_f32=False)
        assert np.array_equal(enc, want)
        assert rr.chunk_info(cx, cy, cz)[0] == 0
    rr.close()


def test_incremental_invalidation_matches_reference(ref, port):
    """Random edits through GameMap::Set -> shadow MarkUpdate -> Update -> radiosity Invalidate: the dirty
    boxes of all 4096 chunks and the bitmap must match after every frame."""
    solid, color = workloads.synthetic_map(3)
    m = ref.map_from_arrays(solid, color)
    rr = oracles.RefRadiosity(ref, m, 1)
    bm = port.shadow_bitmap(solid, color)
    assert np.array_equal(bm, rr.shadow_bitmap())
    # bring the reference to the clean steady state, mirror it on the port side
    assert rr.update_all(8) == 4096 and rr.num_dirty() == 0
    state = np.zeros((4096, 7), np.int32)
    state[:, 4:] = 15
    assert np.array_equal(rr.chunk_state()[:, 0], state[:, 0])
    words = np.zeros(512 * 16, np.uint32)
    rng = np.random.default_rng(99)
    for frame in range(6):
        for _ in range(40):
            x, y = int(rng.integers(0, 512)), int(rng.integers(0, 512))
            col = int(solid[x, y])
            top = (col & -col).bit_length() - 1
            if rng.random() < 0.5 and top < 62:  # dig the top voxel
                z, s = top, False
            else:  # build one above it
                z, s = max(top - 1, 0), True
            c = 0x64000000 | int(rng.integers(0, 1 << 24))
            if s:
                solid[x, y] |= np.uint64(1) << np.uint64(z)
                color[x, y, z] = c
            else:
                solid[x, y] &= ~(np.uint64(1) << np.uint64(z))
            ref.map_set(m, x, y, z, s, c)
            rr.map_changed(x, y, z)
            port.shadow_mark(words, x, y, z)
        rr.shadow_update()
        changed = port.shadow_update(solid, color, bm, words, state)
        assert changed > 0 and not words.any()
        assert np.array_equal(bm, rr.shadow_bitmap())
        want = rr.chunk_state()
        assert np.array_equal(state[:, 0], want[:, 0])
        dirty = want[:, 0] == 1
        assert dirty.any() and np.array_equal(state[dirty], want[dirty])
        if frame % 2 == 1:  # drain: recompute the dirty boxes on both sides and compare the words
            ids = np.nonzero(dirty)[0][:6]
            for i in ids:
                cx, cy, cz = (i // 4) % 32, (i // 4) // 32, i % 4
                lo = (cx * 16 + state[i, 1], cy * 16 + state[i, 2], cz * 16 + state[i, 3])
                hi = (cx * 16 + state[i, 4], cy * 16 + state[i, 5], cz * 16 + state[i, 6])
                _, enc = port.radiosity_box(bm, lo, hi, 1, want_f32=False)
                rr.update_chunk(cx, cy, cz)
                got = rr.read_chunk(cx, cy, cz)[state[i, 3]:state[i, 6] + 1, state[i, 2]:state[i, 5] + 1,
                                                state[i, 1]:state[i, 4] + 1]
                assert np.array_equal(enc, got)
                state[i, 0] = 0
    rr.close()
    ref.map_free(m)
