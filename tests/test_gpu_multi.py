"""GPU, 2 ranks over NCCL (skipped with fewer than 2 GPUs): edit broadcast + view/ray shards through the product
Context on every rank; the union of the shards must equal the single-GPU result and replicas must agree."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import openspades_b200 as osb
from openspades_b200 import sharding, workloads

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


def test_two_rank_nccl_matches_single_gpu(tmp_path):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "w")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"), "--backend", "nccl",
           "--out", out, "--rays", "200000", "--views", "6"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-3000:]
    r = [np.load(f"{out}.rank{k}.npz") for k in range(2)]
    assert len(set(r[0]["digests"].tolist() + r[1]["digests"].tolist())) == 1
    assert int(r[0]["n_gathered"]) == 6
    solid, color = workloads.synthetic_map(11)
    scenes = workloads.random_scenes(6, solid, 64, 64, 68.0, seed=5)
    with osb.Context(0, osb.NUM_IEEE) as ctx:
        ctx.map_upload(solid, color)
        for frame in range(3):
            rng = np.random.default_rng(1000 + frame)
            n = 50
            e = sharding.make_edits(rng.integers(0, 512, n), rng.integers(0, 512, n), rng.integers(1, 62, n),
                                    rng.integers(0, 2, n), 0x64000000 | rng.integers(0, 1 << 24, n))
            sharding.apply_edits_arrays(solid, color, e, ctx)
            ctx.map_flush(solid, color)
        v0, d = workloads.random_rays(200000, seed=77)
        want = ctx.castray2(v0, d, 128).view(np.uint8).reshape(-1, 40)
        fb, _ = ctx.render(scenes, 64, 64, (0.8, 1.0, 1.0), 128.0)
    assert np.array_equal(np.concatenate([r[0]["hits"], r[1]["hits"]]), want)
    assert np.array_equal(np.concatenate([r[0]["fb"], r[1]["fb"]]), fb)
