"""osb200_castray2_static: the CastRay2 batch without any allocation inside the call (static device scratch).

No counterpart in the reference; added for callers that cannot call cudaMalloc.
Collected last on purpose.  The new entry point is compared, through the C ABI, with the
existing osb200_castray2 on the same inputs: bit-exact, like every CastRay2 test here -- on two calls in a row, the second
one smaller than the first so that anything left in the shared scratch by the first call would show."""
import numpy as np
import pytest

import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    m = VoxelMap.load_file(workloads.TITLE_VXL)
    with osb.Context(0, osb.NUM_IEEE) as c:
        c.map_upload(m.solid.copy(), m.color.copy())
        yield c


def _rays(n, seed):
    v0, d = workloads.random_rays(n, seed=seed)
    ev0, ed = workloads.edge_rays()  # zero / NaN / infinite directions, origins outside the map
    k = min(len(ev0), n)
    v0[:k], d[:k] = ev0[:k], ed[:k]
    return v0, d


def test_two_calls_in_a_row_equal_the_allocating_entry_point(ctx):
    nmax = ctx.CASTRAY2_STATIC_MAX_RAYS
    for n, seed, steps in ((nmax, 11, 256), (20001, 12, 64)):
        v0, d = _rays(n, seed)
        want = ctx.castray2(v0, d, steps)
        got = ctx.castray2_static(v0, d, steps, out=np.full(n, 0xAB, np.uint8).repeat(40).view(osb.RAYHIT_DTYPE))
        assert got.tobytes() == want.tobytes(), n
        assert 0.05 < want["hit"].mean() < 1.0


def test_sizes_beyond_the_static_scratch_are_rejected_and_empty_batches_are_fine(ctx):
    n = ctx.CASTRAY2_STATIC_MAX_RAYS + 1
    v0, d = workloads.random_rays(n, seed=3)
    with pytest.raises(osb.OsbError, match="OSB200_CASTRAY2_STATIC_MAX_RAYS"):
        ctx.castray2_static(v0, d, 16)
    assert len(ctx.castray2_static(v0[:0], d[:0], 16)) == 0
    # the failed call left nothing behind: the next one is still right
    want = ctx.castray2(v0[:777], d[:777], 16)
    assert ctx.castray2_static(v0[:777], d[:777], 16).tobytes() == want.tobytes()


def test_a_second_context_on_the_device_shares_the_scratch(ctx):
    """the scratch address is per device, not per context; calls one after the other are fine"""
    m = VoxelMap.load_file(workloads.TITLE_VXL)
    v0, d = _rays(5000, 21)
    with osb.Context(0, osb.NUM_IEEE) as other:
        other.map_upload(m.solid.copy(), m.color.copy())
        a = other.castray2_static(v0, d, 128)
    b = ctx.castray2_static(v0, d, 128)
    assert a.tobytes() == b.tobytes() == ctx.castray2(v0, d, 128).tobytes()
