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
