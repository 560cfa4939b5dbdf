"""Worker for tests/test_sharding_gloo.py (and, with --backend nccl, tests/test_gpu_multi.py): one process per rank.
Rank 0 invents the frame's edits; broadcast_edits ships them; every rank applies them to its replica, then works
on its shard_range of the rays.  On CPU the compute stand-in is the oracle port (test infrastructure); on GPU the
product Context.  Writes a .npz per rank for the parent to compare."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="gloo")
    ap.add_argument("--out", required=True)
    ap.add_argument("--rays", type=int, default=20000)
    ap.add_argument("--views", type=int, default=6)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from openspades_b200 import sharding, workloads
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = None
    if a.backend == "nccl":
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        dist.init_process_group("nccl", device_id=dev)
    else:
        dist.init_process_group("gloo")
    solid, color = workloads.synthetic_map(11)
    ctx = None
    if a.backend == "nccl":
        import openspades_b200 as osb
        ctx = osb.Context(device=local, numerics=osb.NUM_IEEE)
        ctx.map_upload(solid, color)
    v0, d = workloads.random_rays(a.rays, seed=77)
    scenes = workloads.random_scenes(a.views, solid, 64, 64, 68.0, seed=5)
    frames = []
    for frame in range(3):
        edits = None
        if rank == 0:
            rng = np.random.default_rng(1000 + frame)
            n = 50
            edits = sharding.make_edits(rng.integers(0, 512, n), rng.integers(0, 512, n), rng.integers(1, 62, n),
                                        rng.integers(0, 2, n), 0x64000000 | rng.integers(0, 1 << 24, n))
        edits = sharding.broadcast_edits(edits, src=0, device=dev)
        sharding.apply_edits_arrays(solid, color, edits, ctx)
        if ctx is not None:
            ctx.map_flush(solid, color)
        frames.append(len(edits))
    lo, hi = sharding.shard_range(a.rays, rank, world)
    vlo, vhi = sharding.shard_range(a.views, rank, world)
    if ctx is not None:
        hits = ctx.castray2(v0[lo:hi], d[lo:hi], 128)
        fb, dp = ctx.render(scenes[vlo:vhi], 64, 64, (0.8, 1.0, 1.0), 128.0) if vhi > vlo else (np.zeros((0, 64, 64), np.uint32),) * 2
        gathered = sharding.gather_frames(torch.from_numpy(fb.astype(np.int64)).to(dev)) if (a.views % world == 0) else None
        n_gathered = 0 if gathered is None else sum(int(t.shape[0]) for t in gathered)
    else:
        import oracles
        port = oracles.Port()
        hits = port.castray2(solid, v0[lo:hi], d[lo:hi], 128)
        fb = np.stack([port.render(solid, color, scenes[i], 64, 64, (0.8, 1.0, 1.0), 128.0, 0, 4, 1)[0]
                       for i in range(vlo, vhi)]) if vhi > vlo else np.zeros((0, 64, 64), np.uint32)
        n_gathered = 0
    digests = [None] * world
    dist.all_gather_object(digests, sharding.map_digest(solid, color))
    np.savez(f"{a.out}.rank{rank}.npz", hits=hits.view(np.uint8).reshape(-1, 40), fb=fb, lo=lo, hi=hi, vlo=vlo, vhi=vhi,
             digests=np.array(digests), frames=np.array(frames), n_gathered=n_gathered)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
