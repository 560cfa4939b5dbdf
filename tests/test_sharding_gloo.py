"""CPU, world_size 2, gloo: the N>1 host logic -- shard ranges, edit broadcast, replica
consistency, and that the union of the shards equals the single-process result."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import oracles
from openspades_b200 import sharding, workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 4096, 1 << 24, 4097):
        for world in (1, 2, 3, 4, 8):
            r = [sharding.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def test_apply_edits_follows_gamemap_set_semantics():
    solid, color = workloads.synthetic_map(11)
    s0, c0 = solid.copy(), color.copy()
    z_air = 0
    e = sharding.make_edits([3, 3, 3], [4, 4, 4], [z_air, z_air, z_air], [1, 1, 0], [0x64010203, 0x64010203, 0x64999999])
    assert sharding.apply_edits_arrays(solid, color, e[:1]) == 1          # air -> solid
    assert sharding.apply_edits_arrays(solid, color, e[1:2]) == 0         # same again: nothing changes, no notification
    assert sharding.apply_edits_arrays(solid, color, e[2:]) == 1          # removal ignores the colour argument
    assert np.array_equal(solid, s0) and int(color[3, 4, z_air]) == 0x64010203
    color[3, 4, z_air] = c0[3, 4, z_air]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_gloo_broadcast_and_shards(tmp_path):
    out = str(tmp_path / "w")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "dist_worker.py"), "--out", out]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    r = [np.load(f"{out}.rank{k}.npz") for k in range(2)]
    # replicas identical on every rank after three broadcast frames of 50 edits
    assert len(set(r[0]["digests"].tolist() + r[1]["digests"].tolist())) == 1
    assert r[0]["frames"].tolist() == [50, 50, 50] and r[1]["frames"].tolist() == [50, 50, 50]
    # single-process replay of the same edits
    solid, color = workloads.synthetic_map(11)
    for frame in range(3):
        rng = np.random.default_rng(1000 + frame)
        n = 50
        e = sharding.make_edits(rng.integers(0, 512, n), rng.integers(0, 512, n), rng.integers(1, 62, n),
                                rng.integers(0, 2, n), 0x64000000 | rng.integers(0, 1 << 24, n))
        sharding.apply_edits_arrays(solid, color, e)
    assert sharding.map_digest(solid, color) == str(r[0]["digests"][0])
    port = oracles.Port()
    v0, d = workloads.random_rays(20000, seed=77)
    want = port.castray2(solid, v0, d, 128).view(np.uint8).reshape(-1, 40)
    got = np.concatenate([r[0]["hits"], r[1]["hits"]])
    assert (int(r[0]["lo"]), int(r[0]["hi"]), int(r[1]["lo"]), int(r[1]["hi"])) == (0, 10000, 10000, 20000)
    assert np.array_equal(got, want)
    scenes = workloads.random_scenes(6, workloads.synthetic_map(11)[0], 64, 64, 68.0, seed=5)  # as the workers: pre-edit map
    fbs = np.concatenate([r[0]["fb"], r[1]["fb"]])
    assert fbs.shape[0] == 6
    for i in (0, 3, 5):
        fb, _ = port.render(solid, color, scenes[i], 64, 64, (0.8, 1.0, 1.0), 128.0, 0, 4, 1)
        assert np.array_equal(fbs[i], fb)
