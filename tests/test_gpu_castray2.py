"""GPU parity: osb200_castray2 (through the C ABI) vs the oracle restatement and, when it travelled to
this box, the unmodified reference (oracle/_ref).  Bit-exact on every field of RayCastResult."""
import numpy as np
import pytest

import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = osb.Context(device=0, numerics=osb.NUM_IEEE)
    yield c
    c.close()


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


@pytest.fixture(scope="module")
def title_map():
    return VoxelMap.load_file(oracles.TITLE_VXL)


def _same(a, b):
    if a.tobytes() == b.tobytes():
        return True
    for f in ("hit", "startSolid", "hitPos", "hitBlock", "normal"):
        av, bv = a[f].view(np.uint8 if a[f].dtype == np.uint8 else np.uint32), b[f].view(
            np.uint8 if b[f].dtype == np.uint8 else np.uint32)
        bad = np.nonzero(np.any((av != bv).reshape(len(a), -1), axis=1))[0]
        if len(bad):
            print(f"field {f}: {len(bad)} mismatches, first {bad[:5]}: {a[f][bad[:3]]} vs {b[f][bad[:3]]}")
    return False


def test_map_roundtrip(ctx, title_map):
    ctx.map_upload(title_map.solid, title_map.color)
    assert np.array_equal(ctx.map_read_solid(), title_map.solid)


@pytest.mark.parametrize("max_steps", [256, 64, 12])
def test_castray2_title_bit_exact(ctx, port, title_map, max_steps):
    ctx.map_upload(title_map.solid, title_map.color)
    v0, d = workloads.random_rays(1 << 20, seed=1234)
    got = ctx.castray2(v0, d, max_steps)
    want = port.castray2(title_map.solid.copy(), v0, d, max_steps, threads=8)
    assert _same(got, want)
    if oracles.ref_available():
        ref = oracles.Ref("sse2")
        m = ref.map_from_arrays(title_map.solid.copy(), title_map.color.copy())
        n = 200000
        assert _same(got[:n], ref.castray2(m, v0[:n], d[:n], max_steps, threads=8))
        ref.map_free(m)


def test_castray2_edges_and_ragged(ctx, port, title_map):
    ctx.map_upload(title_map.solid, title_map.color)
    v0, d = workloads.edge_rays()  # includes zero, underflowing, infinite and NaN directions
    for n in (0, 1, 31, 33, 255, 257, len(v0)):
        got = ctx.castray2(v0[:n], d[:n], 300)
        want = port.castray2(title_map.solid.copy(), v0[:n], d[:n], 300)
        assert _same(got, want), n


def test_castray2_pageable_and_pinned_paths_agree(ctx, port, title_map):
    """osb200_castray2 moves pageable caller buffers through a pinned ring with a few host threads and pinned ones
    directly; both must return the same records (ragged sizes: not a multiple of either chunk size)."""
    import torch
    ctx.map_upload(title_map.solid, title_map.color)
    n = (1 << 20) + 12345
    v0, d = workloads.random_rays(n, seed=77)
    ev0, ed = workloads.edge_rays()
    v0[1000:1000 + len(ev0)], d[1000:1000 + len(ed)] = ev0, ed
    pageable = ctx.castray2(v0, d, 200)
    pv0, pd = torch.from_numpy(v0).pin_memory(), torch.from_numpy(d).pin_memory()
    pout = torch.zeros(n * 40, dtype=torch.uint8).pin_memory()
    ctx.castray2(pv0, pd, 200, out=pout)
    pinned = pout.numpy().view(pageable.dtype)
    assert pageable.tobytes() == pinned.tobytes()
    k = 1 << 18
    lo = n - k
    assert _same(pageable[lo:], port.castray2(title_map.solid.copy(), v0[lo:], d[lo:], 200, threads=8))
    assert _same(pageable[:k], port.castray2(title_map.solid.copy(), v0[:k], d[:k], 200, threads=8))


def test_castray2_synthetic_and_edit(ctx, port):
    solid, color = workloads.synthetic_map(7)
    ctx.map_upload(solid, color)
    v0, d = workloads.random_rays(300000, seed=5)
    assert _same(ctx.castray2(v0, d, 256), port.castray2(solid, v0, d, 256, threads=8))
    # incremental edits: dirty columns only
    rng = np.random.default_rng(3)
    for _ in range(500):
        x, y, z = int(rng.integers(0, 512)), int(rng.integers(0, 512)), int(rng.integers(0, 63))
        solid[x, y] ^= np.uint64(1) << np.uint64(z)
        color[x, y, z] = 0x64000000 | int(rng.integers(0, 1 << 24))
        ctx.map_mark_dirty(x, y, z)
    n = ctx.map_flush(solid, color)
    assert 0 < n <= 500
    assert np.array_equal(ctx.map_read_solid(), solid)
    assert _same(ctx.castray2(v0, d, 256), port.castray2(solid, v0, d, 256, threads=8))
    assert ctx.map_flush(solid, color) == 0


def test_castray2_16_mi_rays_bit_exact(ctx, port, title_map):
    """north_star verbatim: GameMap::CastRay2 bit-exact on 16 M rays -- 2^24 rays, seed 1234, maxSteps 256, every field of
    RayCastResult against the oracle on all host cores (and the first 2^20 against the unmodified reference)."""
    import os
    ctx.map_upload(title_map.solid, title_map.color)
    n = 1 << 24
    v0, d = workloads.random_rays(n, seed=1234)
    got = ctx.castray2(v0, d, 256)
    threads = max(1, len(os.sched_getaffinity(0)))
    solid = title_map.solid.copy()
    chunk = 1 << 22
    for s in range(0, n, chunk):
        want = port.castray2(solid, v0[s:s + chunk], d[s:s + chunk], 256, threads=threads)
        assert _same(got[s:s + chunk], want), s
    if oracles.ref_available():
        ref = oracles.Ref("sse2")
        m = ref.map_from_arrays(solid, title_map.color.copy())
        k = 1 << 20
        assert _same(got[:k], ref.castray2(m, v0[:k], d[:k], 256, threads=threads))
        ref.map_free(m)
