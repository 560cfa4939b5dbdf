"""Maps for the VXL load / save parity cases: arrays in GameMap's [x][y] / [x][y][z] order."""
import numpy as np


def synthetic_arrays(seed, caves=True):
    """Random terrain with overhangs, floating slabs, holes down to the water floor, full and almost-empty columns, and
    distinct colours everywhere (so that a wrong colour offset shows).  z = 63 is always solid (water floor)."""
    rng = np.random.default_rng(seed)
    xs, ys = np.meshgrid(np.arange(512), np.arange(512), indexing="ij")
    hgt = (40 + 14 * np.sin(xs / 37.0 + seed) * np.cos(ys / 23.0) + rng.integers(-2, 3, (512, 512))).astype(np.int64)
    hgt = np.clip(hgt, 1, 62)
    solid = np.zeros((512, 512), np.uint64)
    full = np.uint64(0xFFFFFFFFFFFFFFFF)
    solid[:] = full << hgt.astype(np.uint64)
    if caves:
        for _ in range(4):  # carve slabs of air below the surface and add floating slabs above it
            z0 = rng.integers(2, 60, (512, 512)).astype(np.uint64)
            th = rng.integers(1, 6, (512, 512)).astype(np.uint64)
            band = ((np.uint64(1) << th) - np.uint64(1)) << z0
            pick = rng.random((512, 512))
            solid = np.where(pick < 0.25, solid & ~band, solid)
            solid = np.where((pick >= 0.25) & (pick < 0.4), solid | band, solid)
    solid[5, 7] = full                      # completely solid column
    solid[6, 7] = np.uint64(1) << np.uint64(63)  # only the water floor
    solid[0, :] |= np.uint64(1)             # voxels at z = 0 on the map edge
    solid[511, 511] = np.uint64(0x8000000000000001)
    solid |= np.uint64(1) << np.uint64(63)
    color = rng.integers(0, 1 << 24, (512, 512, 64), dtype=np.uint32) | np.uint32(100 << 24)
    return solid, color
