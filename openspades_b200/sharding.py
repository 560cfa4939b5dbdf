"""Multi-GPU plumbing for the voxel hot path: one process per GPU, the map replicated,
independent units (views, rays) split into contiguous ranges per rank, and NO collective on the
data path.  The only exchange step is the control path of map edits: rank 0's per-frame edit list is broadcast
(NCCL over NVLink on a GPU box, gloo in the CPU tests) and every rank applies it to its own host map, which fires
the same GameMapChanged -> dirty-column -> osb200_map_flush sequence everywhere, so replicas stay bit-identical.

The reference has nothing distributed; the edit record mirrors the arguments of
GameMap::Set (Sources/Client/GameMap.h:104-135).
"""
import numpy as np

# one GameMap::Set call: 12 bytes
EDIT_DTYPE = np.dtype([("x", "<i2"), ("y", "<i2"), ("z", "<i2"), ("solid", "<u2"), ("color", "<u4")])
assert EDIT_DTYPE.itemsize == 12


def shard_range(n, rank, world):
    """Contiguous share [lo, hi) of n independent units for `rank` of `world` (balanced to within one unit)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return (n * rank) // world, (n * (rank + 1)) // world


def make_edits(xs, ys, zs, solids, colors):
    e = np.zeros(len(xs), EDIT_DTYPE)
    e["x"], e["y"], e["z"], e["solid"], e["color"] = xs, ys, zs, solids, colors
    return e


def broadcast_edits(edits, src=0, device=None, group=None):
    """Broadcasts rank `src`'s edit list to every rank; returns it (EDIT_DTYPE array) on all ranks.
    `device`: torch device of the staging tensors (cuda:<local_rank> under NCCL, None/cpu under gloo)."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return np.ascontiguousarray(edits, EDIT_DTYPE)
    rank = dist.get_rank(group)
    dev = torch.device("cpu") if device is None else device
    cnt = torch.zeros(1, dtype=torch.int64, device=dev)
    if rank == src:
        edits = np.ascontiguousarray(edits, EDIT_DTYPE)
        cnt[0] = len(edits)
    dist.broadcast(cnt, src=src, group=group)
    n = int(cnt.item())
    if n == 0:
        return np.zeros(0, EDIT_DTYPE)
    if rank == src:
        payload = torch.from_numpy(edits.view(np.uint8).reshape(-1).copy()).to(dev)
    else:
        payload = torch.empty(n * EDIT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    dist.broadcast(payload, src=src, group=group)
    return payload.cpu().numpy().view(EDIT_DTYPE).copy()


def apply_edits(vmap, edits):
    """Applies edits to a host VoxelMap through Set() (listeners -- an attached B200Renderer -- fire)."""
    for e in edits:
        vmap.set(int(e["x"]), int(e["y"]), int(e["z"]), bool(e["solid"]), int(e["color"]))


def apply_edits_arrays(solid, color, edits, ctx=None):
    """Same on raw arrays (solid[x][y] uint64, color[x][y][z] uint32) with GameMap::Set's semantics; when a
    Context is given every changed voxel is marked dirty on it (call ctx.map_flush(solid, color) afterwards).
    Returns the number of edits that changed something."""
    changed = 0
    for e in edits:
        x, y, z, s, c = int(e["x"]), int(e["y"]), int(e["z"]), bool(e["solid"]), int(e["color"])
        bit = np.uint64(1) << np.uint64(z)
        was = bool(solid[x, y] & bit)
        ch = False
        if was != s:
            solid[x, y] = (solid[x, y] | bit) if s else (solid[x, y] & ~bit)
            ch = True
        if s and int(color[x, y, z]) != c:
            color[x, y, z] = c
            ch = True
        if ch:
            changed += 1
            if ctx is not None:
                ctx.map_mark_dirty(x, y, z)
    return changed


def gather_frames(t, dst=0, group=None):
    """Optional gather of per-rank result tensors (frames, depth, ray hits) to rank `dst`; off the timed data path.
    Only the root receives (grouped send / recv, not an all-gather: at 8 ranks that is an eighth of the traffic).
    Returns the list of per-rank tensors (rank order) on `dst`, None elsewhere.  Every rank's tensor must have the
    same shape and dtype."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return [t]
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    t = t.contiguous()
    if rank == dst:
        out = [t if r == dst else torch.empty_like(t) for r in range(world)]
        ops = [dist.P2POp(dist.irecv, out[r], r, group) for r in range(world) if r != dst]
    else:
        out = None
        ops = [dist.P2POp(dist.isend, t, dst, group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    return out


def map_digest(solid, color):
    """Order-independent-free, cheap digest of a map replica for cross-rank consistency checks."""
    import hashlib
    h = hashlib.sha1()
    h.update(np.ascontiguousarray(solid).tobytes())
    h.update(np.ascontiguousarray(color).tobytes())
    return h.hexdigest()
