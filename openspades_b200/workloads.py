"""Seeded synthetic inputs for tests and bench.py: maps, rays, camera poses.

Product-side helper: no oracle code is used here.  Everything is numpy on the host; the arrays are
what the C ABI takes (solid uint64[512][512] as [x][y], color uint32[512][512][64] as [x][y][z]).
"""
import math
import os

import numpy as np

from .binding import SCENE_DTYPE

TITLE_VXL = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "maps",
                         "Title.vxl")


def _value_noise(rng, n, cells):
    """Periodic bilinear value noise on an n x n grid with `cells` lattice cells per side."""
    g = rng.random((cells, cells)).astype(np.float64)
    t = np.arange(n) * (cells / n)
    i0 = np.floor(t).astype(int) % cells
    i1 = (i0 + 1) % cells
    f = t - np.floor(t)
    f = f * f * (3 - 2 * f)
    a = g[np.ix_(i0, i0)]
    b = g[np.ix_(i1, i0)]
    c = g[np.ix_(i0, i1)]
    d = g[np.ix_(i1, i1)]
    fx = f[:, None]
    fy = f[None, :]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def synthetic_map(seed=7):
    """Height field (surface z in 8..56) + tileable 3-D caves + floating slabs + hashed colours.

    Periodic in x and y (the reference wraps coordinates with & 511)."""
    rng = np.random.default_rng(seed)
    h = (_value_noise(rng, 512, 4) * 0.55 + _value_noise(rng, 512, 16) * 0.3 + _value_noise(rng, 512, 64) * 0.15)
    top = (8 + h * 48).astype(np.int64)  # first solid z
    z = np.arange(64)[None, None, :]
    vox = z >= top[:, :, None]
    # caves: carve where two tileable noises agree
    c1 = _value_noise(rng, 512, 32)
    c2 = _value_noise(rng, 512, 8)
    cave_z = (20 + c2 * 30).astype(np.int64)
    cave = (np.abs(z - cave_z[:, :, None]) < 3) & (c1[:, :, None] > 0.62)
    vox &= ~cave
    # floating slabs / overhangs
    slab = (_value_noise(rng, 512, 24) > 0.8)
    slab_z = (4 + _value_noise(rng, 512, 6) * 10).astype(np.int64)
    vox |= slab[:, :, None] & (z >= slab_z[:, :, None]) & (z < slab_z[:, :, None] + 2)
    vox[:, :, 63] = True  # water floor is always solid
    weights = (np.uint64(1) << np.arange(64, dtype=np.uint64))
    solid = (vox.astype(np.uint64) * weights[None, None, :]).sum(axis=2, dtype=np.uint64)
    # colours: deterministic hash of position, health byte 100 (some damaged blocks with health < 100)
    xs, ys, zs = np.meshgrid(np.arange(512, dtype=np.uint32), np.arange(512, dtype=np.uint32),
                             np.arange(64, dtype=np.uint32), indexing="ij")
    hsh = (xs * np.uint32(73856093)) ^ (ys * np.uint32(19349663)) ^ (zs * np.uint32(83492791)) ^ np.uint32(seed)
    hsh ^= hsh >> np.uint32(13)
    hsh *= np.uint32(0x5bd1e995)
    hsh ^= hsh >> np.uint32(15)
    base_r = (60 + (zs * 2)).astype(np.uint32)
    r = (base_r + (hsh & 63)) & 255
    g = (90 + ((hsh >> 8) & 127)) & 255
    b = (40 + ((hsh >> 16) & 63)) & 255
    health = np.where((hsh >> 24) % 17 == 0, 50, 100).astype(np.uint32)
    color = (r | (g << 8) | (b << 16) | (health << 24)).astype(np.uint32)
    return np.ascontiguousarray(solid), np.ascontiguousarray(color)


def random_rays(n, seed=1234, zmax=62.0):
    """origins uniform in the map volume (not rejected when inside solid),
    directions uniform on the sphere, un-normalised by a factor in [0.1, 10)."""
    rng = np.random.default_rng(seed)
    v0 = np.empty((n, 3), np.float32)
    v0[:, 0] = rng.random(n, np.float32) * 512
    v0[:, 1] = rng.random(n, np.float32) * 512
    v0[:, 2] = rng.random(n, np.float32) * zmax
    zz = rng.random(n) * 2 - 1
    phi = rng.random(n) * 2 * math.pi
    rr = np.sqrt(np.maximum(0, 1 - zz * zz))
    s = 0.1 + rng.random(n) * 9.9
    d = np.stack([rr * np.cos(phi) * s, rr * np.sin(phi) * s, zz * s], axis=1).astype(np.float32)
    return v0, d


def edge_rays():
    """Axis-aligned directions, zero components, origins on integer boundaries, rays leaving through
    z < 0 and z >= 64, x/y wrap, tiny and huge directions, zero / underflowing / infinite / NaN directions."""
    v0, d = [], []
    origins = [(10.0, 20.0, 5.0), (0.0, 0.0, 0.0), (511.999, 511.999, 1.5), (256.5, 256.5, 0.25), (100.0, 200.5, 30.0),
               (-3.5, 700.25, 10.5), (511.5, 0.5, -2.5), (12.25, 13.75, 62.9), (300.0, 300.0, 63.5), (5.5, 5.5, 70.0)]
    dirs = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1), (1, 1, 0), (1, -1, 0), (0, 1, 1),
            (1, 0, -1), (1, 1, 1), (-1, -1, -1), (1e-3, 1, 0), (1, 1e-6, 1e-6), (1e-20, 1e-20, 1), (3e5, -2e5, 1e4),
            (-0.0, 1, 0), (0.5, 0.25, -0.125), (1, 2, 3), (-3, 2, -1),
            # no direction at all (never moves), a length that underflows to 0, infinite / NaN components
            (0, 0, 0), (1e-30, -1e-30, 1e-30), (float("inf"), 1, 0), (float("nan"), 1, 0), (0, float("nan"), 1),
            (0, 1, float("-inf")), (1, float("nan"), float("nan"))]
    for o in origins:
        for dd in dirs:
            v0.append(o)
            d.append(dd)
    return np.array(v0, np.float32), np.array(d, np.float32)


def look_at_scene(eye, at, up, fov_y_deg, w, h, z_near=0.1, z_far=222.0):
    """Camera basis as the title screen builds it (Resources/Scripts/Gui/MainScreen/MainScreenUI.as:100-128):
    dir = norm(at-eye), side = norm(dir x up), up' = -(dir x side); fovX from the aspect ratio."""
    eye = np.asarray(eye, np.float32)
    d = np.asarray(at, np.float32) - eye
    d = d / np.float32(np.linalg.norm(d))
    side = np.cross(d, np.asarray(up, np.float32)).astype(np.float32)
    side = side / np.float32(np.linalg.norm(side))
    upp = (-np.cross(d, side)).astype(np.float32)
    fov_y = math.radians(fov_y_deg)
    fov_x = 2 * math.atan(math.tan(fov_y / 2) * w / h)
    sc = np.zeros((), SCENE_DTYPE)
    sc["origin"] = eye
    sc["axis"][0] = side
    sc["axis"][1] = upp
    sc["axis"][2] = d
    sc["fovX"] = fov_x
    sc["fovY"] = fov_y
    sc["zNear"] = z_near
    sc["zFar"] = z_far
    return sc


def title_scene(w, h, fov_y_deg=30.0):
    """the title-screen camera frozen at cameraX = 256."""
    return look_at_scene((256.0, 256.0, 12.0), (256.1, 257.0, 12.5), (0.0, 0.0, -1.0), fov_y_deg, w, h)


def random_scenes(n, solid, w, h, fov_y_deg=68.0, seed=4321, max_pitch_deg=60.0):
    """eye over a random column, 2..20 blocks above its top solid voxel (never inside
    solid), yaw uniform, pitch uniform in +-max_pitch (covers the yawRange = pi and the pole / 2pi case)."""
    rng = np.random.default_rng(seed)
    out = np.zeros(n, SCENE_DTYPE)
    k = 0
    while k < n:
        x = rng.random() * 512
        y = rng.random() * 512
        word = int(solid[int(x) & 511, int(y) & 511])
        top = (word & -word).bit_length() - 1 if word else 63
        zf = top - (2 + rng.random() * 18)
        iz = math.floor(zf)
        if iz >= 0 and (word >> iz) & 1:
            continue
        yaw = rng.random() * 2 * math.pi
        pitch = math.radians((rng.random() * 2 - 1) * max_pitch_deg)
        d = (math.cos(yaw) * math.cos(pitch), math.sin(yaw) * math.cos(pitch), math.sin(pitch))
        eye = (x, y, zf)
        at = (eye[0] + d[0], eye[1] + d[1], eye[2] + d[2])
        out[k] = look_at_scene(eye, at, (0.0, 0.0, -1.0), fov_y_deg, w, h)
        k += 1
    return out


def blob_model(seed, w=24, h=24, d=40):
    """Player-sized voxel model for the in-game frame workload: an ellipsoid with a few holes, ~8 % emissive and
    ~8 % colour-0 (customColor) voxels.  Returns the arrays osb200_model_create takes (VoxelModel layout,
    Core/VoxelModel.h:68-106): (w, h, d, origin, solid uint64[h][w], colors uint32[h][w][d])."""
    rng = np.random.default_rng(seed)
    x, y, z = np.meshgrid(np.arange(w), np.arange(h), np.arange(d), indexing="xy")  # [h][w][d] with "xy"
    cx, cy, cz = (w - 1) / 2, (h - 1) / 2, (d - 1) / 2
    r = ((x - cx) / (w / 2)) ** 2 + ((y - cy) / (h / 2)) ** 2 + ((z - cz) / (d / 2)) ** 2
    filled = (r < 1.0) & (rng.random(r.shape) < 0.95)
    colors = rng.integers(1, 1 << 24, size=r.shape, dtype=np.uint32)
    u = rng.random(r.shape)
    colors[u < 0.08] = 0
    colors[(u >= 0.08) & (u < 0.16)] |= np.uint32(1 << 24)
    colors[~filled] = 0
    solid = (filled.astype(np.uint64) << np.arange(d, dtype=np.uint64)[None, None, :]).sum(axis=2).astype(np.uint64)
    origin = np.array([-cx - 0.5, -cy - 0.5, -cz - 0.5], np.float32)
    return w, h, d, origin, solid, colors


def game_frame(scene, handles, n_instances, n_lights, seed=99):
    """In-game frame population in front of `scene`'s camera: n_instances voxel models (handles cycled) at 0.1 block
    per voxel, 4..70 blocks away inside the view cone, random yaw; n_lights point lights of radius 4..16 near them.
    Returns (MODEL_INSTANCE_DTYPE[n_instances], LIGHT_DTYPE[n_lights])."""
    from .binding import LIGHT_DTYPE, MODEL_INSTANCE_DTYPE
    rng = np.random.default_rng(seed)
    sc = np.asarray(scene).reshape(-1)[0]
    o = np.asarray(sc["origin"], np.float64)
    right, up, front = (np.asarray(sc["axis"][i], np.float64) for i in range(3))
    tx, ty = math.tan(float(sc["fovX"]) / 2), math.tan(float(sc["fovY"]) / 2)
    inst = np.zeros(n_instances, MODEL_INSTANCE_DTYPE)
    spots = []
    for i in range(n_instances):
        dist = 4.0 + 66.0 * rng.random() ** 2
        pos = o + front * dist + right * (rng.uniform(-0.9, 0.9) * tx * dist) + up * (rng.uniform(-0.9, 0.9) * ty * dist)
        spots.append(pos)
        ang = rng.uniform(0, 2 * math.pi)
        c, s = math.cos(ang), math.sin(ang)
        m = np.zeros(16, np.float32)
        m[0:3] = (0.1 * c, 0.1 * s, 0)
        m[4:7] = (-0.1 * s, 0.1 * c, 0)
        m[8:11] = (0, 0, 0.1)
        m[12:15] = pos
        m[15] = 1
        inst[i]["model"] = handles[i % len(handles)]
        inst[i]["matrix"] = m
        inst[i]["customColor"] = rng.random(3)
    lights = np.zeros(n_lights, LIGHT_DTYPE)
    for i in range(n_lights):
        base = spots[i % len(spots)] if spots else o + front * 10
        lights[i]["origin"] = base + rng.normal(size=3) * 2.0
        lights[i]["radius"] = rng.uniform(4, 16)
        lights[i]["color"] = rng.uniform(0.3, 1.0, 3)
    return inst, lights
