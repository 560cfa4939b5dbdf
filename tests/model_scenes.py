"""Shared builders for the model / dynamic-light parity cases (tests + golden generator): voxel models, instance
matrices and light sets placed in front of a scene's camera."""
import math
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KV6 = [os.path.join(HERE, "golden", "models", n) for n in ("BlockCursorLine.kv6", "Sight1.kv6")]


def synthetic_model(seed, w=13, h=9, d=17):
    """Random blobby voxel model with holes, emissive voxels (material 1) and colour-0 voxels (-> customColor).
    Returns (w, h, d, origin, solid[h][w] uint64, colors[h][w][d] uint32)."""
    rng = np.random.default_rng(seed)
    solid = np.zeros((h, w), np.uint64)
    colors = np.zeros((h, w, d), np.uint32)
    cx, cy, cz = (w - 1) / 2, (h - 1) / 2, (d - 1) / 2
    for y in range(h):
        for x in range(w):
            bits = 0
            for z in range(d):
                r = ((x - cx) / (w / 2)) ** 2 + ((y - cy) / (h / 2)) ** 2 + ((z - cz) / (d / 2)) ** 2
                if r < 1.0 and rng.random() < 0.93:
                    bits |= 1 << z
                    c = int(rng.integers(1, 1 << 24))
                    u = rng.random()
                    if u < 0.08:
                        c = 0                      # replaced by customColor
                    elif u < 0.16:
                        c |= 1 << 24               # MaterialType::Emissive
                    colors[y, x, z] = c
            solid[y, x] = np.uint64(bits)
    origin = np.array([-cx - 0.5, -cy - 0.5, -cz - 0.5], np.float32)
    return w, h, d, origin, solid, colors


def _rot(axis, ang):
    axis = np.asarray(axis, np.float64)
    axis /= np.linalg.norm(axis)
    x, y, z = axis
    c, s = math.cos(ang), math.sin(ang)
    return np.array([[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                     [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                     [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]])


def matrix(pos, scale, axis=(0, 0, 1), ang=0.0):
    """Column-major Matrix4::m: three basis vectors (rotation * per-axis scale) and the origin."""
    R = _rot(axis, ang) * np.asarray(scale, np.float64)[None, :]
    m = np.zeros(16, np.float32)
    for c in range(3):
        m[c * 4:c * 4 + 3] = R[:, c]
    m[12:15] = pos
    m[15] = 1
    return m


def placements(scene, n_models, seed):
    """[(model index, matrix16, customColor)] spread in front of, beside, very near and behind the camera; the same
    transform appears twice (equal depths: the earlier splat must win)."""
    rng = np.random.default_rng(seed)
    o = np.asarray(scene["origin"], np.float64).reshape(3)
    right, up, front = (np.asarray(scene["axis"], np.float64).reshape(3, 3)[i] for i in range(3))
    out = []
    spots = [(6, 0, 0, 0.25), (10, 2.5, 1, 0.4), (3, -0.8, -0.4, 0.12), (1.2, 0.2, -0.2, 0.05), (25, -6, 2, 0.8),
             (40, 10, -4, 1.0), (-5, 0, 0, 0.3), (8, 30, 0, 0.3), (0.6, 0.0, 0.0, 0.1)]
    for k, (df, dr, du, sc) in enumerate(spots):
        pos = o + front * df + right * dr + up * du
        s = (sc, sc * (1 + 0.5 * (k % 3 == 1)), sc * (1 - 0.3 * (k % 2)))
        m = matrix(pos, s, rng.normal(size=3), float(rng.random() * 6.28))
        cc = rng.random(3).astype(np.float32)
        out.append((k % n_models, m, cc))
        if k == 0:
            out.append(((k + 1) % n_models, m.copy(), rng.random(3).astype(np.float32)))
            out.append((k % n_models, m.copy(), cc))
    return out


def lights(scene, seed):
    """(origin, radius, color) rows: camera inside the radius, ahead, off to the side, behind the view plane, beyond
    the fog cull, overlapping pairs, an over-bright colour."""
    rng = np.random.default_rng(seed)
    o = np.asarray(scene["origin"], np.float64).reshape(3)
    right, up, front = (np.asarray(scene["axis"], np.float64).reshape(3, 3)[i] for i in range(3))
    rows = [(o + front * 1.0, 6.0, (1.0, 0.8, 0.5)), (o + front * 12 + right * 2, 8.0, (0.2, 0.9, 1.0)),
            (o + front * 9 - right * 6 + up * 1, 5.0, (2.0, 0.1, 0.1)), (o - front * 10, 4.0, (1, 1, 1)),
            (o + front * 400, 20.0, (1, 1, 1)), (o + front * 30 - up * 3, 25.0, (0.3, 0.3, 0.9)),
            (o + front * 7 + up * 6, 3.0, (0.5, 1.0, 0.2)), (o + front * 5 + right * 20, 4.0, (1, 0, 1))]
    for _ in range(4):
        rows.append((o + front * rng.uniform(2, 40) + right * rng.uniform(-10, 10) + up * rng.uniform(-5, 5),
                     float(rng.uniform(1, 12)), tuple(rng.random(3))))
    return rows
