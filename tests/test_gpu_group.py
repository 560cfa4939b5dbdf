"""The multi-GPU group of the C ABI (osb200_group_*, csrc/group.cu): one context per GPU inside one process, the map
replicated (PCIe once + ncclBroadcast), edits broadcast as dirty-column records, views and rays sharded by contiguous
range, optional gather of the frames to one device.  Everything must be bit-identical to a single-device context.
A one-device group exercises the code path on the 1-GPU test box; the 2-device cases need 2 GPUs (gpurun --gpus 2)."""
import numpy as np
import pytest

import openspades_b200 as osb
from openspades_b200 import sharding, workloads

pytestmark = pytest.mark.gpu
FOG = (0.8, 1.0, 1.0)


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _edit_frames(solid, color, apply_to, frames=3, n=60):
    for frame in range(frames):
        rng = np.random.default_rng(500 + frame)
        e = sharding.make_edits(rng.integers(0, 512, n), rng.integers(0, 512, n), rng.integers(1, 62, n),
                                rng.integers(0, 2, n), 0x64000000 | rng.integers(0, 1 << 24, n))
        apply_to(e)


@pytest.mark.parametrize("ndev", [1, 2])
def test_group_matches_single_context(ndev):
    if _ngpu() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    solid, color = workloads.synthetic_map(11)
    s1, c1 = solid.copy(), color.copy()
    s2, c2 = solid.copy(), color.copy()
    w = h = 128
    scenes = workloads.random_scenes(37, solid, w, h, 68.0, seed=5)  # 37: shards of unequal size
    v0, d = workloads.random_rays(300001, seed=77)
    with osb.Context(0, osb.NUM_SSE2) as ctx, osb.Group(list(range(ndev)), osb.NUM_SSE2) as grp:
        assert grp.size == ndev
        ctx.map_upload(s1, c1)
        grp.map_upload(s2, c2)
        fb0, dp0 = ctx.render(scenes, w, h, FOG, 128.0)
        fb1, dp1 = grp.render(scenes, w, h, FOG, 128.0)
        assert np.array_equal(fb0, fb1) and np.array_equal(dp0.view(np.uint32), dp1.view(np.uint32))
        assert grp.castray2(v0, d, 128).tobytes() == ctx.castray2(v0, d, 128).tobytes()

        # edits: marked on the group, flushed once (PCIe) + broadcast (NVLink); every replica must follow
        def to_ctx(e):
            sharding.apply_edits_arrays(s1, c1, e, ctx)
            ctx.map_flush(s1, c1)

        class _G:  # apply_edits_arrays only needs map_mark_dirty
            def map_mark_dirty(self, x, y, z):
                grp.map_mark_dirty(x, y, z)

        flushed = []

        def to_grp(e):
            sharding.apply_edits_arrays(s2, c2, e, _G())
            flushed.append(grp.map_flush(s2, c2))
        _edit_frames(solid, color, to_ctx)
        _edit_frames(solid, color, to_grp)
        assert all(f > 0 for f in flushed)
        fb0, dp0 = ctx.render(scenes, w, h, FOG, 128.0)
        fb1, dp1 = grp.render(scenes, w, h, FOG, 128.0)
        assert np.array_equal(fb0, fb1) and np.array_equal(dp0.view(np.uint32), dp1.view(np.uint32))
        assert grp.castray2(v0, d, 128).tobytes() == ctx.castray2(v0, d, 128).tobytes()

        # gather of the frames to the last device of the group
        import torch
        root = ndev - 1
        g_fb = torch.zeros((len(scenes), h, w), dtype=torch.int32, device=f"cuda:{root}")
        g_dp = torch.zeros((len(scenes), h, w), dtype=torch.float32, device=f"cuda:{root}")
        grp.render_gather(scenes, w, h, FOG, 128.0, root, g_fb, g_dp)
        torch.cuda.synchronize(root)
        assert np.array_equal(g_fb.cpu().numpy().view(np.uint32), fb0)
        assert np.array_equal(g_dp.cpu().numpy().view(np.uint32), dp0.view(np.uint32))


def test_group_4096_views_two_gpus_one_call():
    """4096 views on 2 GPUs through one call; bits equal the 1-GPU result"""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    from openspades_b200.hostlib import VoxelMap
    m = VoxelMap.load_file(workloads.TITLE_VXL)
    solid, color = m.solid.copy(), m.color.copy()
    S = 256
    scenes = workloads.random_scenes(4096, solid, S, S, 68.0, seed=4321)
    with osb.Group([0, 1], osb.NUM_SSE2) as grp:
        grp.map_upload(solid, color)
        fb1, _ = grp.render(scenes, S, S, FOG, 128.0, want_depth=False)
    with osb.Context(0, osb.NUM_SSE2) as ctx:
        ctx.map_upload(solid, color)
        fb0, _ = ctx.render(scenes, S, S, FOG, 128.0, want_depth=False)
    assert np.array_equal(fb0, fb1)
