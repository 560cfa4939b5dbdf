"""Pins oracle/orc_links.c (GameMapWrapper::Rebuild / AddBlock / RemoveBlocks, SWFlatMapRenderer::GeneratePixel) against the
unmodified reference compiled in place.  Floating blocks are compared in the reference's order, link maps as bitmaps."""
import numpy as np
import pytest

import links_cases
import oracles


@pytest.fixture(scope="module")
def port():
    return oracles.Port()


def run_script_ref(ref, solid, color, ops):
    """World::ApplyBlockActions order: AddBlock / RemoveBlocks, then the floating cells are cleared with GameMap::Set."""
    m = ref.map_from_arrays(solid, color)
    w = ref.wrapper_create(m)
    outs = [ref.wrapper_linked(w)]
    for op in ops:
        if op[0] == "add":
            if not (int(ref.map_arrays(m)[0][op[1], op[2]]) >> op[3]) & 1:
                ref.wrapper_add_block(w, op[1], op[2], op[3], op[4])
        else:
            s = ref.map_arrays(m)[0]
            cells = [c for c in op[1] if (int(s[c[0], c[1]]) >> c[2]) & 1]
            fl = ref.wrapper_remove_blocks(w, cells) if cells else np.zeros((0, 3), np.int16)
            for (x, y, z) in fl:
                ref.map_set(m, int(x), int(y), int(z), False, 0)
            outs.append(fl)
    outs.append(ref.wrapper_linked(w))
    outs.append(ref.map_arrays(m)[0])
    ref.wrapper_free(w)
    ref.map_free(m)
    return outs


def run_script_port(port, solid, color, ops):
    solid, color = solid.copy(), color.copy()
    w = port.links_create(solid)
    outs = [port.links_linked(w)]
    for op in ops:
        if op[0] == "add":
            if not (int(solid[op[1], op[2]]) >> op[3]) & 1:
                port.links_add_block(w, solid, color, op[1], op[2], op[3], op[4])
        else:
            cells = [c for c in op[1] if (int(solid[c[0], c[1]]) >> c[2]) & 1]
            fl = port.links_remove_blocks(w, solid, cells) if cells else np.zeros((0, 3), np.int16)
            for (x, y, z) in fl:
                solid[x, y] &= ~(np.uint64(1) << np.uint64(z))
            outs.append(fl)
    outs.append(port.links_linked(w))
    outs.append(solid)
    port.links_free(w)
    return outs


def test_links_script_matches_reference(port):
    if not oracles.ref_available("ieee"):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("ieee")
    solid, color = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    ops = links_cases.script(solid, seed=1)
    want = run_script_ref(ref, solid, color, ops)
    got = run_script_port(port, solid, color, ops)
    assert len(want) == len(got)
    fell = 0
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
        if a.ndim == 2 and a.shape[1] == 3:
            fell += len(a)
    assert fell > 40  # the towers, arms and bridge did fall


def test_minimap_matches_reference_flat_map_renderer(port):
    if not oracles.ref_available("ieee"):
        pytest.skip("oracle/_ref not present")
    ref = oracles.Ref("ieee")
    solid, color = port.vxl_load(open(oracles.TITLE_VXL, "rb").read())
    m = ref.map_from_arrays(solid, color)
    rr = oracles.RefRenderer(ref, 64, 64, -1, 1)
    rr.set_map(m)
    want = rr.minimap()
    assert np.array_equal(port.minimap(solid, color), want)
    rr.set_map(None)
    rr.close()
    ref.map_free(m)


def test_clusterize_blocks_port_matches_reference_and_definition(port):
    """orc_clusterize_blocks against World::ClusterizeBlocks itself (the unmodified static function of Client/World.cpp
    compiled into oracle/_ref) and against the definition of a connected component."""
    ref = oracles.Ref("ieee") if oracles.ref_available("ieee") else None
    for name, cells in links_cases.cluster_cases().items():
        k, lab = port.clusterize_blocks(cells)
        got = links_cases.canonical_partition(cells, lab)
        assert len(got) == k, name
        assert got == links_cases.brute_partition(cells), name
        if ref is not None:
            rk, rlab = ref.clusterize_blocks(cells)
            assert (rlab >= 0).all(), name
            assert rk == k and links_cases.canonical_partition(cells, rlab) == got, name
    assert port.clusterize_blocks(np.zeros((0, 3), np.int16))[0] == 0
