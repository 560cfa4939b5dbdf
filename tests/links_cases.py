"""Edit scripts for the floating-block parity cases (GameMapWrapper::AddBlock / RemoveBlocks as World::ApplyBlockActions
drives them): a list of ("add", x, y, z, colour) and ("remove", [(x, y, z), ...]) operations."""
import numpy as np


def top_of(solid, x, y):
    """z of the highest solid voxel of a column (smallest z)."""
    m = int(solid[x, y])
    return (m & -m).bit_length() - 1


def script(solid, seed=0, n_random=150):
    """Towers with arms, bridges between towers, a free-floating cluster, mass digs; returns the op list.  `solid` is
    only read (to find the surface); positions are spread over the map incl. the map edge."""
    rng = np.random.default_rng(seed)
    ops = []
    col = lambda: int(rng.integers(0, 1 << 24)) | (100 << 24)  # noqa: E731
    sites = [(100, 100), (300, 200), (10, 400), (0, 0), (511, 300), (256, 511), (400, 40)]
    for (x, y) in sites:
        z0 = top_of(solid, x, y)
        if z0 < 14:
            continue
        # tower of 8 on the surface, an arm of 5 along +x (or -x at the edge), a hanging bit under the arm
        tower = [(x, y, z0 - 1 - k) for k in range(8)]
        sx = 1 if x < 500 else -1
        arm = [(x + sx * k, y, z0 - 8) for k in range(1, 6)]
        hang = [(x + sx * 5, y, z0 - 7), (x + sx * 5, y, z0 - 6)]
        for c in tower + arm + hang:
            ops.append(("add",) + c + (col(),))
        ops.append(("remove", [arm[1]]))            # cuts the arm: 3 arm cells + 2 hanging cells fall
        ops.append(("remove", [tower[2], tower[5]]))  # two cuts in one call: everything above the lower cut falls
    # bridge between two towers: stays up until the second support goes
    x, y = 200, 300
    za, zb = top_of(solid, x, y), top_of(solid, x + 6, y)
    zt = min(za, zb) - 6
    if zt > 2:
        ta = [(x, y, z) for z in range(za - 1, zt - 1, -1)]
        tb = [(x + 6, y, z) for z in range(zb - 1, zt - 1, -1)]
        br = [(x + k, y, zt) for k in range(1, 6)]
        for c in ta + tb + br:
            ops.append(("add",) + c + (col(),))
        ops.append(("remove", [ta[0]]))   # tower a is cut at its foot: it hangs from the bridge, nothing falls
        ops.append(("remove", [tb[0]]))   # now both are cut: towers + bridge fall
    # a cluster built in mid-air (never linked), then one of its cells is removed: the rest is reported
    x, y = 150, 350
    zf = max(top_of(solid, x, y) - 12, 1)
    fl = [(x, y, zf), (x + 1, y, zf), (x + 1, y + 1, zf), (x + 1, y + 1, zf - 1) if zf > 1 else (x + 2, y + 1, zf)]
    for c in fl:
        ops.append(("add",) + c + (col(),))
    ops.append(("remove", [fl[1]]))
    # random digs of surface voxels (usually nothing falls) and undercuts one below the surface
    for _ in range(n_random):
        x, y = (int(v) for v in rng.integers(0, 512, 2))
        z = top_of(solid, x, y)
        if z >= 62:
            continue
        cells = [(x, y, z)] if rng.random() < 0.5 or z + 1 >= 62 else [(x, y, z + 1)]
        ops.append(("remove", cells))
    # one big batch: a 12 x 12 trench two voxels deep
    x0, y0 = 330, 330
    cells = []
    for dx in range(12):
        for dy in range(12):
            z = top_of(solid, x0 + dx, y0 + dy)
            for dz in range(2):
                if z + dz < 62:
                    cells.append((x0 + dx, y0 + dy, z + dz))
    ops.append(("remove", cells))
    return ops
