"""GPU parity of the floating-block detection (osb200_links_*) and the minimap (osb200_minimap) through the C ABI against
the C restatement of GameMapWrapper / SWFlatMapRenderer::GeneratePixel and, when present, the unmodified reference.
Floating blocks are compared as sets (the GPU returns them sorted, the reference in its walk order); the linked bitmap
and the map after every step bit for bit."""
import numpy as np
import pytest

import links_cases
import oracles
import vxl_cases
import openspades_b200 as osb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


def _sorted(cells):
    cells = np.asarray(cells, np.int16).reshape(-1, 3)
    return cells[np.lexsort((cells[:, 2], cells[:, 1], cells[:, 0]))]


def _run(port, solid, color, ops, ref=None):
    """World::ApplyBlockActions on three implementations in lock-step."""
    ps, pc = solid.copy(), color.copy()      # port's map
    gs, gc = solid.copy(), color.copy()      # host truth of the GPU context
    pw = port.links_create(ps)
    rm = rw = None
    if ref is not None:
        rm = ref.map_from_arrays(solid, color)
        rw = ref.wrapper_create(rm)
    fell = 0
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(gs, gc)
        rounds = ctx.links_rebuild()
        assert rounds >= 1
        assert np.array_equal(ctx.links_read(), port.links_linked(pw))
        if rw is not None:
            assert np.array_equal(ctx.links_read(), ref.wrapper_linked(rw))
        for op in ops:
            if op[0] == "add":
                _, x, y, z, col = op
                if (int(gs[x, y]) >> z) & 1:
                    continue
                port.links_add_block(pw, ps, pc, x, y, z, col)
                if rw is not None:
                    ref.wrapper_add_block(rw, x, y, z, col)
                gs[x, y] |= np.uint64(1) << np.uint64(z)
                gc[x, y, z] = col
                ctx.map_mark_dirty(x, y, z)
                ctx.map_flush(gs, gc)
                ctx.links_blocks_added()
            else:
                cells = [c for c in op[1] if (int(gs[c[0], c[1]]) >> c[2]) & 1]
                if not cells:
                    continue
                want = port.links_remove_blocks(pw, ps, cells)
                for (x, y, z) in cells:
                    gs[x, y] &= ~(np.uint64(1) << np.uint64(z))
                    ctx.map_mark_dirty(x, y, z)
                ctx.map_flush(gs, gc)
                got = ctx.links_remove_blocks(cells)
                assert np.array_equal(got, _sorted(want)), op
                if rw is not None:
                    assert np.array_equal(got, _sorted(ref.wrapper_remove_blocks(rw, cells)))
                fell += len(got)
                for (x, y, z) in got:  # the floating blocks fall: cleared with plain GameMap::Set
                    for s in (ps, gs):
                        s[x, y] &= ~(np.uint64(1) << np.uint64(z))
                    if rm is not None:
                        ref.map_set(rm, int(x), int(y), int(z), False, 0)
                    ctx.map_mark_dirty(int(x), int(y), int(z))
                ctx.map_flush(gs, gc)
        assert np.array_equal(ctx.links_read(), port.links_linked(pw))
        assert np.array_equal(ctx.map_read_solid(), ps)
    port.links_free(pw)
    if rw is not None:
        ref.wrapper_free(rw)
        ref.map_free(rm)
    return fell


def test_floating_blocks_script_on_title(port):
    solid, color = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    ref = oracles.Ref("ieee") if oracles.ref_available("ieee") else None
    ops = links_cases.script(solid, seed=1)
    assert _run(port, solid, color, ops, ref) > 40


def test_floating_blocks_on_a_cave_map(port):
    """Synthetic map full of caves and pre-existing floating slabs (never linked): Rebuild leaves them unlinked, cutting
    into one reports its cluster, and winding connections need several flood launches."""
    solid, color = vxl_cases.synthetic_arrays(2)
    ops = links_cases.script(solid, seed=5, n_random=60)
    # cut into pre-existing floating slabs: pick unlinked solid voxels from the port's rebuild
    w = port.links_create(solid)
    linked = port.links_linked(w)
    port.links_free(w)
    loose = solid & ~linked
    xs, ys = np.nonzero(loose)
    rng = np.random.default_rng(0)
    for k in rng.choice(len(xs), size=min(12, len(xs)), replace=False):
        x, y = int(xs[k]), int(ys[k])
        m = int(loose[x, y])
        ops.append(("remove", [(x, y, (m & -m).bit_length() - 1)]))
    assert len(xs) > 100
    assert _run(port, solid, color, ops) > 40


def test_minimap_matches_port_and_follows_edits(port):
    solid, color = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        assert np.array_equal(ctx.minimap(), port.minimap(solid, color))
        rng = np.random.default_rng(2)
        for _ in range(300):  # dig the top voxel or put one on top
            x, y = (int(v) for v in rng.integers(0, 512, 2))
            z = links_cases.top_of(solid, x, y)
            if rng.random() < 0.5 and z < 62:
                solid[x, y] &= ~(np.uint64(1) << np.uint64(z))
            elif z > 0:
                solid[x, y] |= np.uint64(1) << np.uint64(z - 1)
                color[x, y, z - 1] = np.uint32(int(rng.integers(0, 1 << 24)) | (100 << 24))
            ctx.map_mark_dirty(x, y, z)
        with pytest.raises(osb.OsbError):
            ctx.minimap()  # unflushed edits
        ctx.map_flush(solid, color)
        assert np.array_equal(ctx.minimap(), port.minimap(solid, color))


def _check_numbering(cells, k, lab):
    """osb200_clusterize_blocks numbers the pieces 0 .. k-1 by their smallest cell (CellPos order)."""
    firsts = {}
    for c, l in zip(np.asarray(cells).tolist(), lab.tolist()):
        firsts[l] = min(firsts.get(l, tuple(c)), tuple(c))
    assert sorted(firsts) == list(range(k))
    assert [firsts[l] for l in range(k)] == sorted(firsts.values())


def test_clusterize_blocks_matches_port_and_reference(port):
    """World::ClusterizeBlocks (World.cpp:154-226): the partition of every case equals the C restatement's and the
    unmodified reference function's; numbering is the documented one; input order does not matter."""
    ref = oracles.Ref("ieee") if oracles.ref_available("ieee") else None
    with osb.Context(0, osb.NUM_IEEE) as ctx:  # needs no map
        assert ctx.clusterize_blocks(np.zeros((0, 3), np.int16))[0] == 0
        for name, cells in links_cases.cluster_cases().items():
            k, lab = ctx.clusterize_blocks(cells)
            got = links_cases.canonical_partition(cells, lab)
            pk, plab = port.clusterize_blocks(cells)
            assert k == pk == len(got), name
            assert got == links_cases.canonical_partition(cells, plab), name
            if ref is not None:
                rk, rlab = ref.clusterize_blocks(cells)
                assert rk == k and got == links_cases.canonical_partition(cells, rlab), name
            _check_numbering(cells, k, lab)
            perm = np.random.default_rng(3).permutation(len(cells))
            k2, lab2 = ctx.clusterize_blocks(cells[perm])
            assert k2 == k and np.array_equal(lab2, lab[perm]), name
        for bad in ([[512, 0, 0]], [[0, -1, 0]], [[0, 0, 64]]):
            with pytest.raises(osb.OsbError):
                ctx.clusterize_blocks(np.array(bad, np.int16))
        assert ctx.launch_count >= 3


def test_clusterize_blocks_on_the_title_terrain(port):
    """Every solid voxel of Title.vxl above z = 48 (hundreds of thousands of cells: hills, buildings, loose pieces)."""
    solid, _ = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    xs, ys = np.nonzero(solid & np.uint64((1 << 48) - 1))
    cells = []
    for z in range(48):
        sel = ((solid[xs, ys] >> np.uint64(z)) & np.uint64(1)).astype(bool)
        cells.append(np.stack([xs[sel], ys[sel], np.full(sel.sum(), z)], 1))
    cells = np.concatenate(cells).astype(np.int16)
    cells = cells[np.random.default_rng(1).permutation(len(cells))]
    assert len(cells) > 100000
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        k, lab = ctx.clusterize_blocks(cells)
    pk, plab = port.clusterize_blocks(cells)
    assert k == pk
    # same partition <=> the pairs (our label, port's label) form a bijection
    pairs = np.unique(np.stack([lab, plab], 1), axis=0)
    assert len(pairs) == k and len(np.unique(pairs[:, 0])) == k and len(np.unique(pairs[:, 1])) == k
    # property at full size: no face-adjacent pair of cells carries two labels
    vol = np.full((512, 512, 48), -1, np.int32)
    vol[cells[:, 0], cells[:, 1], cells[:, 2]] = lab
    for ax in range(3):
        a = np.moveaxis(vol, ax, 0)
        both = (a[1:] >= 0) & (a[:-1] >= 0)
        assert np.array_equal(a[1:][both], a[:-1][both])


def test_wrapper_clusterize_after_remove_blocks(port):
    """World::ApplyBlockActions through the C++ host layer: RemoveBlocks, then ClusterizeBlocks of what fell."""
    from openspades_b200.hostlib import B200Renderer, VoxelMap, VoxelMapWrapper
    solid, color = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    # three towers on a flat piece of ground (one with an L-shaped arm); cutting all three feet in one call drops three
    # separate pieces
    tops = np.array([[links_cases.top_of(solid, x, y) for y in range(512)] for x in range(512)])
    x0 = y0 = None
    for x in range(8, 480, 4):
        for y in range(8, 480, 4):
            win = tops[x - 2:x + 12, y - 2:y + 12]
            if win.min() == win.max() and win.min() >= 12:
                x0, y0 = x, y
                break
        if x0 is not None:
            break
    assert x0 is not None, "no flat 14 x 14 site on the map"
    top = int(tops[x0, y0])
    vm = VoxelMap.from_arrays(solid, color)
    r = B200Renderer(64, 64, numerics=osb.NUM_IEEE)
    r.init()
    r.set_game_map(vm)
    w = VoxelMapWrapper(vm, r)
    w.rebuild()
    posts, want = [], []
    for i, (dx, dy) in enumerate([(0, 0), (6, 0), (0, 6)]):
        x, y = x0 + dx, y0 + dy
        t = links_cases.top_of(solid, x, y)
        col = [(x, y, z) for z in range(t - 1, t - 8, -1)]
        extra = [(x + 1, y, t - 7), (x + 1, y + 1, t - 7)] if i == 2 else []
        for c in col + extra:
            w.add_block(*c, 0x00ff00)
        posts.append(col[0])
        want.append(tuple(sorted(col[1:] + extra)))
    fl = w.remove_blocks(posts)
    pieces = w.clusterize_blocks(fl)
    assert sorted(tuple(map(tuple, p.tolist())) for p in pieces) == sorted(want)
    assert [tuple(p[0]) for p in pieces] == sorted(tuple(p[0]) for p in pieces)  # ordered by smallest cell
    assert top > 8
    w.close()
    r.close()
