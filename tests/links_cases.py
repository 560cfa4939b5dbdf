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
    # voxels added at the three cells Rebuild links through its value-initialised queue (GameMapWrapper.cpp:62): AddBlock
    # has no such path, so in mid-air they stay unlinked; removing one of them then reports nothing else
    quirk = [c for c in ((1, 0, 0), (0, 1, 0), (0, 0, 1)) if top_of(solid, c[0], c[1]) > 3 and top_of(solid, 0, 0) > 3]
    for c in quirk:
        ops.append(("add",) + c + (col(),))
    if quirk:
        ops.append(("remove", [quirk[0]]))
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


def canonical_partition(cells, labels):
    """A labelling as a sorted list of sorted tuples of cells: what World::ClusterizeBlocks returns, without its order."""
    pieces = {}
    for c, l in zip(np.asarray(cells).reshape(-1, 3).tolist(), np.asarray(labels).tolist()):
        pieces.setdefault(l, set()).add(tuple(c))
    return sorted(tuple(sorted(p)) for p in pieces.values())


def cluster_cases(seed=0):
    """Cell sets for World::ClusterizeBlocks: name -> int16 [n, 3]."""
    rng = np.random.default_rng(seed)
    cases = {}
    cases["single"] = [[7, 9, 3]]
    cases["two_apart"] = [[1, 1, 1], [3, 1, 1]]
    cases["diagonal_is_not_connected"] = [[4, 4, 4], [5, 5, 4], [6, 6, 5], [5, 4, 5]]
    cases["corners_of_the_map"] = [[0, 0, 0], [511, 511, 63], [0, 511, 0], [511, 0, 63], [0, 0, 1], [511, 511, 62],
                                   [511, 510, 63], [510, 511, 63]]
    # keys that differ by one in the packed order but are not neighbours: (x, y, 63) and (x, y + 1, 0), (x, 511, z) and (x + 1, 0, z)
    cases["no_wrap_between_columns"] = [[10, 10, 63], [10, 11, 0], [20, 511, 5], [21, 0, 5], [30, 511, 63], [31, 0, 0]]
    line = [[x, 100, 20] for x in range(50, 300)]
    cases["long_line_shuffled"] = rng.permutation(np.array(line)).tolist()
    # a winding path: worst case for label propagation, nothing special for union-find
    snake = []
    for row in range(0, 40, 2):
        xs = range(100, 180) if (row // 2) % 2 == 0 else range(179, 99, -1)
        snake += [[x, 200 + row, 30] for x in xs]
        snake.append([179 if (row // 2) % 2 == 0 else 100, 200 + row + 1, 30])
    cases["snake"] = rng.permutation(np.array(snake)).tolist()
    box = [[x, y, z] for x in range(300, 312) for y in range(300, 310) for z in range(10, 18)]
    shell = [c for c in box if c[0] in (300, 311) or c[1] in (300, 309) or c[2] in (10, 17)]
    cases["hollow_box_with_core"] = shell + [[305, 305, 13], [305, 305, 14], [306, 305, 14]]
    cases["duplicates"] = [[5, 5, 5], [5, 5, 6], [5, 5, 5], [9, 9, 9], [9, 9, 9], [5, 6, 6]]
    pts = rng.integers(0, [24, 24, 16], size=(3000, 3)) + [400, 60, 40]  # ~33 % fill: many pieces of all sizes
    cases["random_cloud"] = np.unique(pts, axis=0)[rng.permutation(len(np.unique(pts, axis=0)))].tolist()
    towers = [[x, y, z] for x in range(0, 512, 37) for y in range(0, 512, 41) for z in range(0, 64)]
    cases["towers_full_height"] = towers
    return {k: np.array(v, np.int16).reshape(-1, 3) for k, v in cases.items()}


def brute_partition(cells):
    """Connected components by definition (pure Python flood over a set), for small cases."""
    todo = {tuple(c) for c in np.asarray(cells).reshape(-1, 3).tolist()}
    out = []
    while todo:
        stack = [todo.pop()]
        piece = []
        while stack:
            c = stack.pop()
            piece.append(c)
            for d in ((1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)):
                nb = (c[0] + d[0], c[1] + d[1], c[2] + d[2])
                if nb in todo:
                    todo.remove(nb)
                    stack.append(nb)
        out.append(tuple(sorted(piece)))
    return sorted(out)
