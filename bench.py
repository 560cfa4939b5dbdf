#!/usr/bin/env python
"""bench.py -- throughput of the OpenSpades voxel hot path on B200 (contract: see the task brief).

Headline workload (BASELINE.json configs[3], the one the north_star's roofline target is quoted on and the
only render workload that shards): a batch of independent 256x256 camera views of the bundled
Title.vxl map, fovY 68 deg, fog (0.8,1,1)/128, V-SSE2 numerics with r_swNumThreads = 4 -- 4096 views PER
GPU (weak scaling: views are sharded by view, map replicated, no data-path collective).
A "step" renders the whole batch once: k_view_prep, k_line_setup, k_line_compact, k_walk_lines (the terrain pass,
SWMapRenderer::BuildLine) and k_gather_fog (RenderFinal + ApplyFog fused) -- one launch of each per step.

  value   Mpixel/s, device-timed, inputs resident in HBM (scene constants upload is part of the call).
  e2e     same metric through the host-pointer C-ABI call osb200_render: host scenes in, frames copied back
          to pinned host memory inside the timed region (depth stays device-resident, as in SWRenderer).
  extra   configs[0] (800x600 title frame), configs[1] (1920x1080 single view), configs[2] (16 Mi CastRay2 rays),
          the in-game frame and configs[4] (radiosity) measured the same way, each with its CPU baseline beside it.

`--impl reference` times the reference's own CPU code (oracle/_ref, else the oracle port) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpixel/s voxel terrain render (batched 256x256 views, terrain + fog)"
UNIT = "Mpixel/s"
FOG = (0.8, 1.0, 1.0)  # Client.cpp:112-113
FOG_DIST = 128.0
BYTES_PER_PIXEL = 8  # 4 B BGRX + 4 B f32 depth written once per pixel
BYTES_PER_RAY = 64   # 24 B in + 40 B RayCastResult out
CPU_VIEWS_PER_WORKER = 256  # CPU sample of both arms: views per worker process per step
STRONG_VIEWS_TOTAL = 4096   # BASELINE.json configs[3]: 4096 views sharded by view across 1/2/4/8 GPUs
FLOP_PER_TAP = 32           # float operations one executed radiosity tap issues (GLRadiosityRenderer.cpp:216-262) now that its
                            # square root and two divisions come from a table (round 1 / early round 2: 42 with those counted as 1 each)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--views", type=int, default=4096, help="views per GPU")
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--rays", type=int, default=1 << 24, help="rays per GPU for the CastRay2 extra")
    ap.add_argument("--no-extra", action="store_true", help="skip the 1080p / CastRay2 extras")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU work per cpu_baseline sample")
    return ap.parse_args()


def load_map():
    """Title.vxl decoded by the product's host VoxelMap (GameMap::Load equivalent)."""
    from openspades_b200 import workloads
    from openspades_b200.hostlib import VoxelMap
    m = VoxelMap.load_file(workloads.TITLE_VXL)
    return m.solid.copy(), m.color.copy()


def load_map_cpu_only():
    """Same arrays without touching the product libraries (reference arm): oracle VXL decoder."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracles
    return oracles.Port().vxl_load(open(oracles.TITLE_VXL, "rb").read())


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, t0, t1):
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.15 and len(r) >= 9]
        if not rows:
            rows = [r for (_, r) in self.rows if len(r) >= 9]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[1]) for r in rows)
        reasons = set()
        for r in rows:
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7),
                              ("sw_power_cap", 8)):
                if r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except (KeyError, ValueError):
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def read_traffic(kernel, workload, views_per_launch):
    """dram bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json): NOT a live
    counter -- a profiler cannot run inside the timed region.  Returns (bytes scaled to this run's launch size, label)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            t = json.load(open(p)).get(workload, {})
            if kernel in t:
                return t[kernel] * views_per_launch / t.get("views_per_launch", views_per_launch), t.get("source")
        except ValueError:
            pass
    return None, None


def cpu_views_sample(args, solid, color, S, per_worker):
    """Reference SWRenderer on the host cores over a bounded sample of the same view distribution."""
    from oracle import cpu_bench
    from openspades_b200 import workloads
    workers = cpu_bench.usable_cores()
    n = workers * per_worker
    scenes = workloads.random_scenes(n, solid, S, S, 68.0, seed=4321)
    return cpu_bench.Pool("views", solid, color, scenes, workers=workers, w=S, h=S, fog=list(FOG),
                          fog_dist=FOG_DIST, sw_threads=1), n


def run_reference(args):
    """--impl reference: the reference's own CPU code, all host cores, same metric / config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import cpu_bench
    S = args.size
    solid, color = load_map_cpu_only()
    # bounded sample: 256 views per worker process per step (~1 s of work per core per step) -- the same sample, pose
    # seed and worker layout as the cpu_baseline leg of the GPU arm, so that the two CPU numbers of one box agree
    pool, n = cpu_views_sample(args, solid, color, S, CPU_VIEWS_PER_WORKER)
    with pool:
        for _ in range(args.warmup):
            pool.step()
        secs = 0.0
        for _ in range(args.steps):
            secs += pool.step()["seconds"]
    ms = secs * 1e3 / args.steps
    value = n * S * S / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32+u8", "data": "synthetic poses on the bundled Title.vxl map",
        "config": {"workload": f"C4 batched {S}x{S} views, fovY 68, fog (0.8,1,1)/128, Title.vxl",
                   "views_per_step": n, "width": S, "height": S},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": pool.workers, "kind": pool.kind,
                         "sample": f"{n} views per step = {CPU_VIEWS_PER_WORKER} per worker process x {pool.workers} processes, each a private "
                                   "SWRenderer (V-SSE2, r_swNumThreads=1), StartScene+EndScene+FrameDone"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_json(line)
    return 0


def parity_views(ctx, scenes, d_fb, d_dp, solid, color, S, n_sample=64, seed=7):
    """Compares a seeded sample of the views the timed steps rendered (frames + depth still on the device) with the
    unmodified reference SWRenderer (oracle/_ref, V-SSE2, r_swNumThreads = 4) or, where it did not travel, the oracle
    port.  Checker only: runs after the timed regions."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracles
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(len(scenes), size=min(n_sample, len(scenes)), replace=False))
    fb = d_fb[idx.tolist()].cpu().numpy().view(np.uint32)
    dp = d_dp[idx.tolist()].cpu().numpy()
    use_ref = oracles.ref_available("sse2") and os.path.exists(os.path.join(oracles.ORACLE_DIR, "_ref", "liboref_sse2.so"))
    same_px = same_dp = within1 = 0
    total = 0
    if use_ref:
        ref = oracles.Ref("sse2")
        m = ref.map_from_arrays(solid, color)
        r = oracles.RefRenderer(ref, S, S, -1, 4)
        r.set_map(m)
        r.set_fog(FOG, FOG_DIST)
    else:
        port = oracles.Port()
    for k, i in enumerate(idx):
        if use_ref:
            rfb, rdp = r.render(scenes[i])
        else:
            rfb, rdp = port.render(solid, color, scenes[i], S, S, FOG, FOG_DIST, 2, 4, 1)
        a, b = fb[k] & 0xFFFFFF, rfb & 0xFFFFFF
        same_px += int((a == b).sum())
        d = np.abs(((a[..., None] >> np.array([0, 8, 16], np.uint32)) & 0xFF).astype(int)
                   - ((b[..., None] >> np.array([0, 8, 16], np.uint32)) & 0xFF).astype(int)).max(axis=-1)
        within1 += int((d <= 1).sum())
        same_dp += int((dp[k].view(np.uint32) == rdp.view(np.uint32)).sum())
        total += a.size
    if use_ref:
        r.set_map(None)
        r.close()
        ref.map_free(m)
    return {"views": int(len(idx)), "against": "oracle/_ref SWRenderer (unmodified reference)" if use_ref else "oracle port",
            "identical": same_px / total, "within_1_lsb": within1 / total, "depth_identical": same_dp / total}


def parity_rays(d_out, v0, d, solid, n_sample=1 << 20):
    """First n_sample rays of the timed CastRay2 batch against the oracle on all host cores: every field bit-exact."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracles
    from oracle import cpu_bench
    n = min(n_sample, len(v0))
    got = d_out[:n].cpu().numpy().view(np.uint8).reshape(n, 40)
    want = oracles.Port().castray2(solid, v0[:n], d[:n], 256, threads=cpu_bench.usable_cores())
    return {"rays": int(n), "bit_exact": bool(got.tobytes() == want.tobytes())}


def run_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import openspades_b200 as osb
    from openspades_b200 import workloads

    S, nv, K, W = args.size, args.views, args.steps, args.warmup
    solid, color = load_map()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- cpu_baseline leg (rank 0, N = 1 only), before the GPU is busy ----
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            pool, n_cpu = cpu_views_sample(args, solid, color, S, CPU_VIEWS_PER_WORKER)
            with pool:
                pool.step()  # warm
                r = pool.step()
            cpu_baseline = {
                "value": n_cpu * S * S / r["seconds"] / 1e6, "unit": UNIT, "cores": pool.workers, "kind": pool.kind,
                "sample": f"{n_cpu} of the same {S}x{S} views ({CPU_VIEWS_PER_WORKER} per worker process x {pool.workers} processes, each a "
                          f"private SWRenderer V-SSE2 r_swNumThreads=1), {r['seconds']:.2f} s wall"}
        except Exception as e:  # the baseline is a reported extra; never let it sink the GPU numbers
            cpu_baseline = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(e)}

    ctx = osb.Context(device=local, numerics=osb.NUM_SSE2, sw_threads=4)
    ctx.map_upload(solid, color)
    ext = torch.cuda.ExternalStream(ctx.stream, device=dev)
    scenes = workloads.random_scenes(nv, solid, S, S, 68.0, seed=4321 + rank)
    d_fb = torch.empty((nv, S, S), dtype=torch.int32, device=dev)
    d_dp = torch.zeros((nv, S, S), dtype=torch.float32, device=dev)
    pixels_per_step = nv * S * S

    def step_device():
        ctx.render_device(scenes, S, S, FOG, FOG_DIST, d_fb, d_dp)

    # ---- device-timed value ----
    for _ in range(W):
        step_device()
    ctx.sync()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record(ext)
    for _ in range(K):
        step_device()
    e1.record(ext)
    e1.synchronize()
    barrier()
    t_wall1 = time.time()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    gpu_launches = ctx.launch_count - launches0
    ms_per_step = ms_total / K
    # per-kernel durations: a second pass of K steps with CUDA events around every launch on the launching stream
    # (the timed region above runs without them).
    ctx.timing_enable(True)
    ctx.timing_read()
    for _ in range(K):
        step_device()
    ctx.sync()
    ktimes = ctx.timing_read()
    ctx.timing_enable(False)
    value = world * pixels_per_step / (ms_per_step * 1e-3) / 1e6

    # ---- what the numerics choice costs: the same step with the other two variants (rank-local, not part of `value`) ----
    numerics_cost = {"V-SSE2": ms_per_step}
    if not args.no_extra:
        for name, num in (("V-IEEE", osb.NUM_IEEE), ("V-SSE", osb.NUM_SSE)):
            with osb.Context(device=local, numerics=num, sw_threads=4) as nc:
                nc.map_upload(solid, color)
                nx = torch.cuda.ExternalStream(nc.stream, device=dev)
                for _ in range(W):
                    nc.render_device(scenes, S, S, FOG, FOG_DIST, d_fb, d_dp)
                nc.sync()
                n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                n0.record(nx)
                for _ in range(K):
                    nc.render_device(scenes, S, S, FOG, FOG_DIST, d_fb, d_dp)
                n1.record(nx)
                n1.synchronize()
                numerics_cost[name] = n0.elapsed_time(n1) / K
        step_device()  # leave the V-SSE2 frames in d_fb / d_dp for the parity sample and the checksum
        ctx.sync()

    # ---- e2e through the host-pointer C-ABI call (H2D scenes, D2H frames + depth every step) ----
    # What the reference-facing call returns is the framebuffer (SWPort bitmap); the depth buffer is internal
    # state of SWRenderer (SWRenderer.h:59) and stays device-resident here too (B200Renderer::EndScene passes
    # host_depth = NULL), so the per-step D2H is 4 B/pixel.
    h_fb = torch.empty((nv, S, S), dtype=torch.int32, pin_memory=True)

    def step_e2e():
        ctx.render(scenes, S, S, FOG, FOG_DIST, want_depth=False, out_bgrx=h_fb)

    for _ in range(max(1, min(W, 2))):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        step_e2e()
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / K
    t_wall2 = time.time()
    e2e = {"value": world * pixels_per_step / (e2e_ms * 1e-3) / 1e6, "unit": UNIT,
           "h2d_bytes_per_step": int(scenes.nbytes), "d2h_bytes_per_step": int(pixels_per_step * 4),
           "ms_per_step": e2e_ms, "note": "host scenes in, BGRX frames out to pinned host memory (depth stays on the "
                                          "device like SWRenderer::depthBuffer); D2H of chunk k overlaps the kernels of chunk k+1"}
    clocks = sampler.summary(t_wall0, t_wall2)
    checksum = int(h_fb.view(-1)[:: 4099].to(torch.int64).sum().item())  # the step's result was read on the host

    # ---- the headline number carries its own correctness: a seeded sample of the rendered views vs the reference ----
    parity = None
    if rank == 0:
        step_device()
        ctx.sync()
        try:
            parity = parity_views(ctx, scenes, d_fb, d_dp, solid, color, S)
        except Exception as e:  # a checker problem must not sink the measurement; it is reported
            parity = {"error": repr(e)}

    # ---- strong scaling: BASELINE.json's config is 4096 views TOTAL, sharded by view over the N GPUs ----
    from openspades_b200 import sharding
    sv = STRONG_VIEWS_TOTAL
    lo, hi = sharding.shard_range(sv, rank, world)
    if world == 1 and nv == sv:
        strong_ms = ms_per_step
    else:
        sc_all = workloads.random_scenes(sv, solid, S, S, 68.0, seed=4321)
        mine = np.ascontiguousarray(sc_all[lo:hi])
        if hi - lo <= nv:
            s_fb, s_dp = d_fb[:hi - lo], d_dp[:hi - lo]
        else:  # --views smaller than this rank's share of the strong-scaling batch: buffers of their own
            s_fb = torch.empty((hi - lo, S, S), dtype=torch.int32, device=dev)
            s_dp = torch.zeros((hi - lo, S, S), dtype=torch.float32, device=dev)
        for _ in range(W):
            ctx.render_device(mine, S, S, FOG, FOG_DIST, s_fb, s_dp)
        ctx.sync()
        barrier()
        e0.record(ext)
        for _ in range(K):
            ctx.render_device(mine, S, S, FOG, FOG_DIST, s_fb, s_dp)
        e1.record(ext)
        e1.synchronize()
        barrier()
        strong_ms = max_over_ranks(e0.elapsed_time(e1)) / K
    t1_equiv = ms_per_step * sv / nv  # one GPU rendering all sv views (the weak run's per-GPU load, scaled if --views differs)
    strong = {"views_total": sv, "views_per_gpu": hi - lo, "ms_per_step": strong_ms,
              "mpixel_s": sv * S * S / (strong_ms * 1e-3) / 1e6,
              "efficiency_vs_n1": (t1_equiv / strong_ms) / world,
              "note": "efficiency = (ms of one GPU rendering all 4096 views, from this run's weak step) / (N x ms of N GPUs "
                      "rendering 4096 / N each); a launch of few hundred views per GPU is dominated by fixed costs (per-call "
                      "host set-up of the views, five launches, the latency of the longest yaw lines)"}

    # ---- the one collective of the design (control path): edit broadcast + dirty-column flush; optional gather to rank 0 ----
    multi_gpu = None
    if world > 1:
        solid_w, color_w = solid.copy(), color.copy()
        with osb.Context(device=local, numerics=osb.NUM_SSE2, sw_threads=4) as ec:
            ec.map_upload(solid_w, color_w)
            rng = np.random.default_rng(99)
            frames, per_frame = 20, 100
            ed_frames = []
            for _ in range(frames):
                xs, ys = rng.integers(0, 512, per_frame), rng.integers(0, 512, per_frame)
                ed = []
                for x, y in zip(xs, ys):
                    col = int(solid[x, y])
                    top = (col & -col).bit_length() - 1
                    ed.append((x, y, max(top - 1, 0), 1, 0x64000000 | int(rng.integers(0, 1 << 24))))
                ed_frames.append(sharding.make_edits(*zip(*ed)))
            barrier()
            t0 = time.perf_counter()
            t_bcast = 0.0
            for f in range(frames):
                tb = time.perf_counter()
                e = sharding.broadcast_edits(ed_frames[f] if rank == 0 else None, 0, dev)
                t_bcast += time.perf_counter() - tb
                sharding.apply_edits_arrays(solid_w, color_w, e, ec)
                ec.map_flush(solid_w, color_w)
            ec.sync()
            barrier()
            edit_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / frames
            bcast_ms = max_over_ranks(t_bcast * 1e3) / frames
            digest = sharding.map_digest(solid_w, color_w)
        dig = [None] * world
        dist.all_gather_object(dig, digest)
        # gather of the finished frames to rank 0: grouped send / recv (not an all_gather: only the root receives)
        reps = 3
        gb = d_fb.numel() * 4 / 1e9
        bufs = [torch.empty_like(d_fb) for _ in range(world - 1)] if rank == 0 else None
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            if rank == 0:
                ops = [dist.P2POp(dist.irecv, bufs[r - 1], r) for r in range(1, world)]
            else:
                ops = [dist.P2POp(dist.isend, d_fb, 0)]
            for req in dist.batch_isend_irecv(ops):
                req.wait()
            torch.cuda.synchronize()
        barrier()
        gather_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / reps
        del bufs
        multi_gpu = {"edit_broadcast": {"edits_per_frame": per_frame, "frames": frames, "ms_per_frame_total": edit_ms,
                                        "ms_per_frame_broadcast": bcast_ms, "replicas_identical": len(set(dig)) == 1,
                                        "what": "rank 0's edit list (12 B per GameMap::Set) broadcast over NCCL, applied to every "
                                                "rank's host map, dirty columns flushed to its GPU"},
                     "gather_to_rank0": {"gb_per_rank": gb, "ms": gather_ms,
                                         "root_ingress_gb_s": gb * (world - 1) / (gather_ms * 1e-3),
                                         "value_with_gather_mpixel_s": world * pixels_per_step / ((ms_per_step + gather_ms) * 1e-3) / 1e6,
                                         "what": "BGRX frames of every rank sent to rank 0 (grouped isend / irecv), after the "
                                                 "render; not part of `value`"}}

    # ---- roofline of the dominant kernel (CUDA events around every launch, on the launching stream) ----
    peak, peak_src = read_peaks()
    dom = max(ktimes, key=lambda k: ktimes[k][1])
    dl, dms = ktimes[dom]
    bytes_per_launch = BYTES_PER_PIXEL * pixels_per_step * K / max(dl, 1)
    achieved = bytes_per_launch / (dms / max(dl, 1) * 1e-3) / 1e9
    traffic, traffic_src = read_traffic(dom, "views256", nv / max(dl / K, 1))
    step_bytes = BYTES_PER_PIXEL * pixels_per_step
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms_per_step * 1e-3) / 1e9,
                         "frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak,
                         "note": "the whole step (all kernels + gaps) against the same 8 B/pixel"},
                "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": dms / max(dl, 1),
                "kernel_timing": "separate pass of the same K steps on the context's stream with CUDA events around "
                                 "every launch (OSB200 kernel timing facility); `value` comes from the untimed-launch pass",
                "kernel_ms_per_step": {k: v[1] / K for k, v in ktimes.items() if v[0]},
                "kernel_launches_per_step": {k: v[0] / K for k, v in ktimes.items() if v[0]}}

    extra = {}
    if not args.no_extra:
        extra = run_extras(args, ctx, ext, solid, color, dev, world, barrier, max_over_ranks, peak)
    sampler.stop()

    if rank == 0 and isinstance(parity, dict) and isinstance(extra.get("castray2", {}).get("parity"), dict):
        parity.update(extra["castray2"]["parity"])
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32+u8", "data": "synthetic poses (seeded) on the bundled Title.vxl map",
            "config": {"workload": f"C4 batched {S}x{S} views, fovY 68, fog (0.8,1,1)/128, Title.vxl, V-SSE2 numerics "
                                   "(r_swNumThreads=4)", "views_per_gpu": nv, "views_total": nv * world, "width": S,
                       "height": S, "parallelism": f"views sharded over {world} GPU(s), map replicated",
                       "l2": "no flush: each step writes 8 B/pixel x %.2f Gpixel = %.2f GB per GPU plus its strip "
                             "buffers, far beyond the 126 MB L2; the 2 MiB map bitmap is L2-resident by design"
                             % (pixels_per_step / 1e9, pixels_per_step * 8 / 1e9)},
            "e2e": e2e, "gpu_launches": int(gpu_launches), "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu_baseline, "parity_sample": parity, "strong": strong, "multi_gpu": multi_gpu,
            "numerics_ms_per_step": numerics_cost, "extra": extra, "result_checksum": checksum,
        }
        emit_json(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def rank_is_zero():
    return int(os.environ.get("RANK", "0")) == 0


def run_extras(args, ctx, ext, solid, color, dev, world, barrier, max_over_ranks, peak):
    """configs[1] (1920x1080 single view; replicas only, rank-local) and configs[2] (CastRay2 batch,
    rays sharded by ray, aggregate over ranks).  Same timing rules: warm-up, events on the context stream."""
    import torch
    from openspades_b200 import workloads
    import openspades_b200 as osb
    out = {}
    rank = int(os.environ.get("RANK", "0"))
    do_cpu = rank == 0 and world == 1 and not args.no_cpu
    if do_cpu:
        from oracle import cpu_bench
    # ---- C1: 800x600, the title-screen camera of the bundled map (fovY 30, fog (0.1,0.1,0.1)/128), the reference's own
    # CPU-runnable case: SWRenderer with its default 4 worker threads beside it ----
    w, h = 800, 600
    sc = workloads.title_scene(w, h, 30.0)
    fb = torch.empty((1, h, w), dtype=torch.int32, device=dev)
    dp = torch.zeros((1, h, w), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    n_it = 20
    for _ in range(3):
        ctx.render_device(sc, w, h, (0.1, 0.1, 0.1), 128.0, fb, dp)
    ctx.sync()
    tot = 0.0
    for _ in range(n_it):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        ctx.render_device(sc, w, h, (0.1, 0.1, 0.1), 128.0, fb, dp)
        e1.record(ext)
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    c1_ms = tot / n_it
    hfb = torch.empty((1, h, w), dtype=torch.int32, pin_memory=True)
    ctx.render(sc, w, h, (0.1, 0.1, 0.1), 128.0, want_depth=False, out_bgrx=hfb)
    t0 = time.perf_counter()
    for _ in range(n_it):
        ctx.render(sc, w, h, (0.1, 0.1, 0.1), 128.0, want_depth=False, out_bgrx=hfb)
    c1_e2e = (time.perf_counter() - t0) * 1e3 / n_it
    out["frame_800x600_title"] = {"workload": "C1 800x600 single view, title-screen camera fovY 30, fog (0.1,0.1,0.1)/128, Title.vxl, "
                                              "V-SSE2 (L2 flushed between iterations)", "ms_per_frame": c1_ms,
                                  "mpixel_s": w * h / c1_ms / 1e3, "e2e_ms_per_frame": c1_e2e, "e2e_mpixel_s": w * h / c1_e2e / 1e3,
                                  "hbm_frac_algorithmic": (BYTES_PER_PIXEL * w * h / (c1_ms * 1e-3) / 1e9) / peak}
    if do_cpu:
        try:
            cms = cpu_bench.single_view_ms(solid, color, sc, w, h, (0.1, 0.1, 0.1), 128.0, threads=4, reps=10)
            out["frame_800x600_title"]["cpu_baseline"] = {"value": w * h / cms / 1e3, "unit": "Mpixel/s", "ms_per_frame": cms,
                                                          "cores": 4, "kind": cpu_bench.kind(),
                                                          "sample": "the same frame, one SWRenderer, r_swNumThreads = 4 (the reference's "
                                                                    "default), best of 10"}
        except Exception as e:
            out["frame_800x600_title"]["cpu_baseline"] = {"value": None, "kind": "unavailable", "sample": repr(e)}
    del fb, dp, hfb, flush
    # ---- 1920x1080 single view, title pose + fog (0.1,0.1,0.1)/128 ----
    w, h = 1920, 1080
    sc = workloads.title_scene(w, h, 68.0)
    fb = torch.empty((1, h, w), dtype=torch.int32, device=dev)
    dp = torch.zeros((1, h, w), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    n_it = 20
    for _ in range(3):
        ctx.render_device(sc, w, h, (0.1, 0.1, 0.1), 128.0, fb, dp)
    ctx.sync()
    tot = 0.0
    for _ in range(n_it):
        flush.fill_(1)  # L2 flush between timed iterations (torch current stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        ctx.render_device(sc, w, h, (0.1, 0.1, 0.1), 128.0, fb, dp)
        e1.record(ext)
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / n_it
    hfb = torch.empty((1, h, w), dtype=torch.int32, pin_memory=True)
    hdp = torch.empty((1, h, w), dtype=torch.float32, pin_memory=True)
    ctx.render(sc, w, h, (0.1, 0.1, 0.1), 128.0, out_bgrx=hfb, out_depth=hdp)
    t0 = time.perf_counter()
    for _ in range(n_it):
        ctx.render(sc, w, h, (0.1, 0.1, 0.1), 128.0, out_bgrx=hfb, out_depth=hdp)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / n_it
    out["frame_1080p"] = {"workload": "C2 1920x1080 single view, title pose fovY 68, fog, Title.vxl (per GPU replica; L2 "
                                      "flushed between iterations)", "ms_per_frame": ms,
                          "mpixel_s": w * h / ms / 1e3, "e2e_ms_per_frame": e2e_ms, "e2e_mpixel_s": w * h / e2e_ms / 1e3,
                          "hbm_frac_algorithmic": (BYTES_PER_PIXEL * w * h / (ms * 1e-3) / 1e9) / peak}
    if do_cpu:
        try:
            cms = cpu_bench.single_view_ms(solid, color, sc, w, h, (0.1, 0.1, 0.1), 128.0, threads=4, reps=5)
            out["frame_1080p"]["cpu_baseline"] = {"value": w * h / cms / 1e3, "unit": "Mpixel/s", "ms_per_frame": cms, "cores": 4,
                                                  "kind": cpu_bench.kind(),
                                                  "sample": "the same frame, one SWRenderer, r_swNumThreads = 4, best of 5"}
        except Exception as e:
            out["frame_1080p"]["cpu_baseline"] = {"value": None, "kind": "unavailable", "sample": repr(e)}
    # ---- the same frame as the game submits it: + 64 voxel models (3 distinct, 24x24x40) + 8 point lights, full EndScene ----
    handles = [ctx.model_create(*workloads.blob_model(11 + k)) for k in range(3)]
    inst, lights = workloads.game_frame(sc, handles, 64, 8)
    ctx.timing_enable(False)
    for _ in range(3):
        ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, to_host=False)
    tot = 0.0
    for _ in range(n_it):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, to_host=False)
        e1.record(ext)
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    g_ms = tot / n_it
    ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, out_bgrx=hfb, out_depth=hdp)
    t0 = time.perf_counter()
    for _ in range(n_it):
        ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, out_bgrx=hfb, out_depth=hdp)
    g_e2e = (time.perf_counter() - t0) * 1e3 / n_it
    ctx.timing_enable(True)
    ctx.timing_read()
    for _ in range(5):
        ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, to_host=False)
    g_kt = {k: v[1] / 5 for k, v in ctx.timing_read().items() if v[0]}
    ctx.timing_enable(False)
    vox = sum(int(ctx.model_render_data(hd, 24 * 24)[0].size - 24 * 24) for hd in handles)
    for hd in handles:
        ctx.model_destroy(hd)
    out["frame_1080p_game"] = {"workload": "1920x1080 title pose fovY 68, full EndScene: terrain + 64 voxel-model instances (3 models of "
                                           "24x24x40, 0.1 block per voxel, 4..70 blocks away) + 8 point lights (radius 4..16) + fog "
                                           "(osb200_render_scene; L2 flushed between iterations)",
                               "ms_per_frame": g_ms, "e2e_ms_per_frame": g_e2e, "fps_device": 1e3 / g_ms, "fps_e2e": 1e3 / g_e2e,
                               "surface_voxels_of_the_3_models": vox, "kernel_ms": g_kt}
    del fb, dp, hfb, hdp, flush

    # ---- the batched views once more on SURVEY.md 8(d)'s synthetic map (height field + caves + floating slabs, hashed
    # colours: more ceilings and overhangs per column than Title.vxl), rank-local ----
    if rank_is_zero():
        S2, nv2 = 256, 1024
        syn_solid, syn_color = workloads.synthetic_map(7)
        with osb.Context(device=dev.index, numerics=osb.NUM_SSE2, sw_threads=4) as sc_ctx:
            sc_ctx.map_upload(syn_solid, syn_color)
            sx = torch.cuda.ExternalStream(sc_ctx.stream, device=dev)
            scenes2 = workloads.random_scenes(nv2, syn_solid, S2, S2, 68.0, seed=99)
            fb2 = torch.empty((nv2, S2, S2), dtype=torch.int32, device=dev)
            dp2 = torch.zeros((nv2, S2, S2), dtype=torch.float32, device=dev)
            for _ in range(3):
                sc_ctx.render_device(scenes2, S2, S2, FOG, FOG_DIST, fb2, dp2)
            sc_ctx.sync()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(sx)
            for _ in range(5):
                sc_ctx.render_device(scenes2, S2, S2, FOG, FOG_DIST, fb2, dp2)
            a1.record(sx)
            a1.synchronize()
            syn_ms = a0.elapsed_time(a1) / 5
            sc_ctx.timing_enable(True)
            sc_ctx.timing_read()
            sc_ctx.render_device(scenes2, S2, S2, FOG, FOG_DIST, fb2, dp2)
            sc_ctx.sync()
            syn_kt = {k: v[1] for k, v in sc_ctx.timing_read().items() if v[0]}
        out["views_synthetic_map"] = {"workload": f"{nv2} views of {S2}x{S2}, fovY 68, fog, workloads.synthetic_map(7) (SURVEY.md 8d), V-SSE2; "
                                                  "1024 x 65536 x 8 B = 537 MB per step > L2",
                                      "ms_per_step": syn_ms, "mpixel_s": nv2 * S2 * S2 / syn_ms / 1e3, "kernel_ms": syn_kt}
        del fb2, dp2, syn_solid, syn_color

    # ---- CastRay2, args.rays rays per GPU, maxSteps 256 ----
    n = args.rays
    rank = int(os.environ.get("RANK", "0"))
    v0, d = workloads.random_rays(n, seed=1234 + rank)
    d_v0 = torch.from_numpy(v0).to(dev)
    d_d = torch.from_numpy(d).to(dev)
    d_out = torch.empty((n, 10), dtype=torch.int32, device=dev)
    for _ in range(2):
        ctx.castray2_device(d_v0, d_d, 256, d_out, n)
    ctx.sync()
    K = 5
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(K):
        ctx.castray2_device(d_v0, d_d, 256, d_out, n)
    e1.record(ext)
    e1.synchronize()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / K
    hits = int((d_out[:, 0] & 1).sum().item())
    # e2e through osb200_castray2 (host buffers in and out; 1 Mi-ray chunks pipelined H2D / kernel / D2H on three streams):
    # from pinned buffers (what a caller that cares about throughput passes) and from ordinary pageable numpy arrays
    ne = n
    p_v0 = torch.from_numpy(v0).pin_memory()
    p_d = torch.from_numpy(d).pin_memory()
    p_out = torch.empty((ne, 10), dtype=torch.int32, pin_memory=True)
    ctx.castray2(p_v0, p_d, 256, out=p_out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        ctx.castray2(p_v0, p_d, 256, out=p_out)
    barrier()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / 3
    e2e_same = bool(torch.equal(p_out, d_out.cpu()))
    pg_out = np.zeros(ne, osb.binding.RAYHIT_DTYPE)
    ctx.castray2(v0, d, 256, out=pg_out)
    t0 = time.perf_counter()
    ctx.castray2(v0, d, 256, out=pg_out)
    pageable_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    del p_v0, p_d, p_out, pg_out
    out["castray2"] = {"workload": f"C3 {n} random rays per GPU, maxSteps 256, Title.vxl (in+out {n * 64 / 1e9:.2f} GB > L2)",
                       "ms_per_batch": ms, "mrays_s": world * n / ms / 1e3, "hit_fraction": hits / n,
                       "hbm_frac_algorithmic": (BYTES_PER_RAY * n / (ms * 1e-3) / 1e9) / peak,
                       "e2e_mrays_s": world * ne / e2e_ms / 1e3, "e2e_rays": ne, "e2e_pageable_mrays_s": world * ne / pageable_ms / 1e3,
                       "e2e_equals_device_result": e2e_same,
                       "e2e_note": "pinned host buffers: 24 B/ray up, 40 B/ray down per step; PCIe-bound on the 40 B/ray direction"}
    if rank == 0:
        try:
            out["castray2"]["parity"] = parity_rays(d_out, v0, d, solid)
        except Exception as e:
            out["castray2"]["parity"] = {"error": repr(e)}
    # strong scaling of C3: 2^24 rays TOTAL split over the ranks
    total_rays = 1 << 24
    if world > 1:
        from openspades_b200 import sharding
        lo, hi = sharding.shard_range(total_rays, rank, world)
        m = min(hi - lo, n)
        for _ in range(2):
            ctx.castray2_device(d_v0, d_d, 256, d_out, m)
        ctx.sync()
        barrier()
        e0.record(ext)
        for _ in range(K):
            ctx.castray2_device(d_v0, d_d, 256, d_out, m)
        e1.record(ext)
        e1.synchronize()
        barrier()
        sms = max_over_ranks(e0.elapsed_time(e1)) / K
        out["castray2"]["strong"] = {"rays_total": total_rays, "rays_per_gpu": m, "ms_per_batch": sms,
                                     "mrays_s": total_rays / sms / 1e3,
                                     "efficiency_vs_n1": (ms * total_rays / n / sms) / world}
    if do_cpu:
        try:
            cores = cpu_bench.usable_cores()
            per = 250_000  # ~0.5 s per core at the reference's ~0.5 Mrays/s/core
            with cpu_bench.Pool("rays", solid, color, (v0[:cores * per], d[:cores * per]), workers=cores, max_steps=256) as pool:
                pool.step()
                r_all = pool.step()
            with cpu_bench.Pool("rays", solid, color, (v0[:per], d[:per]), workers=1, max_steps=256) as pool:
                pool.step()
                r_one = pool.step()
            out["castray2"]["cpu_baseline"] = {"value": r_all["units"] / r_all["seconds"] / 1e6, "unit": "Mrays/s", "cores": cores,
                                               "kind": cpu_bench.kind(), "one_core_mrays_s": r_one["units"] / r_one["seconds"] / 1e6,
                                               "sample": f"{cores * per} of the same rays (GameMap::CastRay2, maxSteps 256), {per} per "
                                                         f"worker process x {cores} processes, {r_all['seconds']:.2f} s wall; and {per} "
                                                         "rays on one core"}
        except Exception as e:
            out["castray2"]["cpu_baseline"] = {"value": None, "kind": "unavailable", "sample": repr(e)}
    del d_v0, d_d, d_out

    # ---- C5: full-map radiosity + incremental edits (GLRadiosityRenderer CPU half) ----
    from openspades_b200 import sharding
    solid2, color2 = load_map()
    with osb.Context(device=dev.index, numerics=osb.NUM_IEEE, radiosity_mode=1) as rc:
        rc.map_upload(solid2, color2)
        rc.shadow_bitmap()
        rc.timing_enable(True)
        rc.timing_read()
        t0 = time.perf_counter()
        n_chunks = rc.radiosity_update(-1)
        full_wall_ms = (time.perf_counter() - t0) * 1e3
        kt = rc.timing_read()
        full_ms = kt["k_radiosity"][1]
        taps_done, taps_all = rc.radiosity_count_taps()
        # 100 frames x 100 random dig/build edits: mark_dirty -> flush -> shadow update -> radiosity(all dirty)
        rng = np.random.default_rng(99)
        frames, per_frame = 100, 100
        t0 = time.perf_counter()
        upd = 0
        for _ in range(frames):
            xs, ys = rng.integers(0, 512, per_frame), rng.integers(0, 512, per_frame)
            ed = []
            for x, y in zip(xs, ys):
                col = int(solid2[x, y])
                top = (col & -col).bit_length() - 1
                if rng.random() < 0.5 and top < 62:
                    ed.append((x, y, top, 0, 0))
                else:
                    ed.append((x, y, max(top - 1, 0), 1, 0x64000000 | int(rng.integers(0, 1 << 24))))
            e = sharding.make_edits(*zip(*ed))
            sharding.apply_edits_arrays(solid2, color2, e, rc)
            rc.map_flush(solid2, color2)
            upd += rc.radiosity_update(-1)
        inc_s = time.perf_counter() - t0
        kt2 = rc.timing_read()
        # ---- ambient half of C5 (GLAmbientShadowRenderer): full map, then 20 frames x 100 edits, drain every frame ----
        rc.timing_read()
        t0 = time.perf_counter()
        ao_chunks = rc.ambient_update(-1)
        rc.sync()
        ao_wall_ms = (time.perf_counter() - t0) * 1e3
        ao_ms = rc.timing_read()["k_ao_chunk"][1]
        ao_frames = 20
        t0 = time.perf_counter()
        ao_upd = 0
        for _ in range(ao_frames):
            xs, ys = rng.integers(0, 512, per_frame), rng.integers(0, 512, per_frame)
            ed = []
            for x, y in zip(xs, ys):
                col = int(solid2[x, y])
                top = (col & -col).bit_length() - 1
                if rng.random() < 0.5 and top < 62:
                    ed.append((x, y, top, 0, 0))
                else:
                    ed.append((x, y, max(top - 1, 0), 1, 0x64000000 | int(rng.integers(0, 1 << 24))))
            e = sharding.make_edits(*zip(*ed))
            sharding.apply_edits_arrays(solid2, color2, e, rc)
            rc.map_flush(solid2, color2)
            ao_upd += rc.ambient_update(-1)
        rc.sync()
        ao_inc_s = time.perf_counter() - t0
        ao_kt = rc.timing_read()
        # ---- GameMap::CastRay batch: 16 Mi unit rays of length 128 (the hit-test form), device resident ----
        import torch
        nr = min(args.rays, 1 << 24)
        rv0, rd = workloads.random_rays(nr, seed=4321)
        rd /= np.linalg.norm(rd, axis=1, keepdims=True)
        t_v0, t_d = torch.from_numpy(rv0).to(dev), torch.from_numpy(rd.astype(np.float32)).to(dev)
        t_out = torch.empty((nr, 4), dtype=torch.int32, device=dev)
        rc.castray_device(t_v0, t_d, 128.0, t_out, nr)
        rc.sync()
        rc.timing_read()
        for _ in range(5):
            rc.castray_device(t_v0, t_d, 128.0, t_out, nr)
        rc.sync()
        cr_ms = rc.timing_read()["k_castray"][1] / 5
        cr_hits = float((t_out[:, 0] != 0).float().mean().item())
    out["ambient"] = {"workload": "C5 ambient half, Title.vxl: all 4096 chunks dirty -> osb200_ambient_update(all) (16 rays x <=16 "
                                  "cell walls per air voxel of each 20^3 window, 2 x 3 in-place blurs), then 20 frames x 100 random "
                                  "block edits, each frame flush + update of every dirty chunk",
                      "full_map_chunks": ao_chunks, "full_map_kernel_ms": ao_ms, "full_map_wall_ms": ao_wall_ms,
                      "full_map_mvoxels_s": 512 * 512 * 64 / ao_ms / 1e3,
                      "window_mrays_s_nominal": 4096 * 8000 * 16 / ao_ms / 1e3,  # rays the reference would cast; voxels out of reach of any solid are skipped exactly
                      "edits_per_s": ao_frames * per_frame / ao_inc_s, "chunk_updates": ao_upd,
                      "incremental_kernel_ms_per_frame": {k: v[1] / ao_frames for k, v in ao_kt.items() if v[0]}}
    out["castray"] = {"workload": f"GameMap::CastRay batch, {nr} unit rays of length 128 per GPU, Title.vxl (in 24 B + out 16 B per ray)",
                      "ms_per_batch": cr_ms, "mrays_s": nr / cr_ms / 1e3, "hit_fraction": cr_hits,
                      "hbm_frac_algorithmic": (40 * nr / (cr_ms * 1e-3) / 1e9) / peak}
    # ---- VXL stream <-> resident map (GameMap::Load / Save): device decode vs host decode + 66 MiB upload ----
    from openspades_b200.hostlib import VoxelMap
    vxl = open(workloads.TITLE_VXL, "rb").read()
    with osb.Context(device=dev.index, numerics=osb.NUM_IEEE) as vc:
        reps = 10
        vc.map_load_vxl(vxl, want_arrays=False)
        vc.timing_enable(True)
        vc.timing_read()
        t0 = time.perf_counter()
        for _ in range(reps):
            vc.map_load_vxl(vxl, want_arrays=False)
        load_ms = (time.perf_counter() - t0) * 1e3 / reps
        load_kt = {k: v[1] / reps for k, v in vc.timing_read().items() if v[0]}
        t0 = time.perf_counter()
        for _ in range(3):
            hs, hc = vc.map_load_vxl(vxl, want_arrays=True)
        load_arrays_ms = (time.perf_counter() - t0) * 1e3 / 3
        vc.timing_read()
        t0 = time.perf_counter()
        for _ in range(reps):
            saved = vc.map_save_vxl()
        save_ms = (time.perf_counter() - t0) * 1e3 / reps
        save_kt = {k: v[1] / (2 * reps) for k, v in vc.timing_read().items() if v[0]}  # each save = length query + emit
        t0 = time.perf_counter()
        for _ in range(3):
            hm = VoxelMap.load(vxl)
            vc.map_upload(hm.solid, hm.color)
        host_ms = (time.perf_counter() - t0) * 1e3 / 3
    out["vxl"] = {"workload": f"Title.vxl ({len(vxl)} B stream, 262144 columns): osb200_map_load_vxl (H2D of the stream, column "
                              "index by pointer doubling, per-column decode into the resident map) and osb200_map_save_vxl, "
                              "against the host decoder (VoxelMap::Load = GameMap::Load) + osb200_map_upload of the 66 MiB arrays",
                  "device_load_ms": load_ms, "device_load_with_host_arrays_ms": load_arrays_ms, "device_save_ms": save_ms,
                  "host_load_plus_upload_ms": host_ms, "saved_bytes": len(saved), "load_kernel_ms": load_kt,
                  "save_kernel_ms_per_call": save_kt, "columns_per_s_load": 262144 / (load_ms * 1e-3)}
    # ---- floating-block detection (GameMapWrapper) + minimap ----
    ls, lc = load_map()
    with osb.Context(device=dev.index, numerics=osb.NUM_IEEE) as lk:
        lk.map_upload(ls, lc)
        lk.links_rebuild()
        t0 = time.perf_counter()
        for _ in range(10):
            rounds = lk.links_rebuild()
        rebuild_ms = (time.perf_counter() - t0) * 1e3 / 10
        rng = np.random.default_rng(5)
        n_calls, per_call, fell = 50, 100, 0
        batches = []
        for _ in range(n_calls):  # 100 distinct columns per call: the voxel under the surface is dug out (undercut)
            cols = rng.choice(512 * 512, per_call, replace=False)
            batches.append([(int(c) >> 9, int(c) & 511) for c in cols])
        lk.timing_enable(True)
        lk.timing_read()
        t0 = time.perf_counter()
        for cols in batches:
            cells = []
            for (x, y) in cols:
                m_ = int(ls[x, y])
                z = (m_ & -m_).bit_length() - 1 + 1
                if z >= 62 or not (m_ >> z) & 1:
                    continue
                ls[x, y] &= ~(np.uint64(1) << np.uint64(z))
                lk.map_mark_dirty(x, y, z)
                cells.append((x, y, z))
            lk.map_flush(ls, lc)
            fl = lk.links_remove_blocks(cells)
            fell += len(fl)
            for (x, y, z) in fl:
                ls[x, y] &= ~(np.uint64(1) << np.uint64(int(z)))
                lk.map_mark_dirty(int(x), int(y), int(z))
        lk.map_flush(ls, lc)
        remove_ms = (time.perf_counter() - t0) * 1e3 / n_calls
        lkt = lk.timing_read()
        lk.timing_read()
        t0 = time.perf_counter()
        for _ in range(20):
            lk.minimap()
        minimap_ms = (time.perf_counter() - t0) * 1e3 / 20
        mm_k = lk.timing_read()["k_minimap"][1] / 20
        # World::ClusterizeBlocks: the solid voxels of Title.vxl above z = 48 as one cell set (hills, buildings, bridges)
        cl_x, cl_y = np.nonzero(ls & np.uint64((1 << 48) - 1))
        cl_cells = []
        for z in range(48):
            sel = ((ls[cl_x, cl_y] >> np.uint64(z)) & np.uint64(1)).astype(bool)
            cl_cells.append(np.stack([cl_x[sel], cl_y[sel], np.full(int(sel.sum()), z)], 1))
        cl_cells = np.ascontiguousarray(np.concatenate(cl_cells).astype(np.int16))
        lk.clusterize_blocks(cl_cells)
        lk.timing_read()
        t0 = time.perf_counter()
        for _ in range(5):
            cl_k, _lab = lk.clusterize_blocks(cl_cells)
        cluster_ms = (time.perf_counter() - t0) * 1e3 / 5
        cluster_k_ms = lk.timing_read()["k_cluster_union"][1] / 5
    out["links"] = {"workload": "GameMapWrapper on Title.vxl: Rebuild (flood from the ground over 16.8 M voxels as 64-bit column "
                                f"masks), then {n_calls} RemoveBlocks calls x {per_call} undercut voxels each (host edit + flush + flood + "
                                "result read-back inside the wall time); minimap = SWFlatMapRenderer image of the whole map",
                    "rebuild_ms": rebuild_ms, "rebuild_flood_launches": rounds, "remove_blocks_ms_per_call": remove_ms,
                    "removed_voxels_per_s": n_calls * per_call / (remove_ms * n_calls * 1e-3), "blocks_fell": fell,
                    "flood_kernel_ms_per_call": lkt["k_links_flood"][1] / n_calls,
                    "flood_launches_per_call": lkt["k_links_flood"][0] / n_calls,
                    "minimap_ms": minimap_ms, "minimap_kernel_ms": mm_k,
                    "clusterize": {"workload": "World::ClusterizeBlocks of every solid voxel of Title.vxl above z = 48 in one call "
                                               "(H2D of the cells, device key sort, union-find, piece numbering, labels D2H inside the wall time)",
                                   "cells": int(len(cl_cells)), "pieces": int(cl_k), "wall_ms": cluster_ms,
                                   "union_kernel_ms": cluster_k_ms,
                                   "mcells_s_kernel": len(cl_cells) / (cluster_k_ms * 1e-3) / 1e6}}
    out["radiosity"] = {"workload": "C5 Title.vxl: all 4096 chunks dirty -> osb200_radiosity_update(all) (16.8 M voxels, <=169 "
                                    "taps each), then 100 frames x 100 random block edits, each frame flush + shadow update + "
                                    "radiosity update of every dirty chunk (host edit generation inside the wall time)",
                        "full_map_chunks": n_chunks, "full_map_kernel_ms": full_ms, "full_map_wall_ms": full_wall_ms,
                        "full_map_mvoxels_s": 512 * 512 * 64 / full_ms / 1e3,
                        "edits_per_s": frames * per_frame / inc_s, "chunk_updates": upd,
                        "incremental_kernel_ms_per_frame": {k: v[1] / frames for k, v in kt2.items() if v[0]}}
    # C5 is bound by FP32 issue, not by HBM: achieved FP32 rate from the EXECUTED taps (those that pass
    # Evaluate's normal cull, counted by a separate kernel) against SMs x 128 lanes x clock; the code is built without FMA
    # contraction (bit parity with the x86 reference), so its ceiling is one operation per lane per clock, half the FMA peak.
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    clk_ghz = 1.965
    flops = taps_done * FLOP_PER_TAP
    achieved_tf = flops / (full_ms * 1e-3) / 1e12
    peak_nofma = sm_count * 128 * clk_ghz / 1e3
    out["radiosity"]["roofline"] = {"bound": "fp32 issue (CUDA cores), no tensor cores: nothing here is a contraction",
                                    "executed_taps": taps_done, "tap_slots": taps_all, "executed_fraction": taps_done / taps_all,
                                    "flop_per_tap": FLOP_PER_TAP, "achieved": achieved_tf, "unit": "TFLOP/s",
                                    "peak_no_fma": peak_nofma, "peak_fma": 2 * peak_nofma, "frac_of_no_fma_peak": achieved_tf / peak_nofma,
                                    "frac_of_fma_peak": achieved_tf / (2 * peak_nofma),
                                    "peak_source": f"{sm_count} SMs x 128 FP32 lanes x {clk_ghz} GHz (x 2 with FMA)"}
    if do_cpu:
        try:
            cores = cpu_bench.usable_cores()
            s1, v1 = cpu_bench.radiosity_sample(solid, color, 32, 1)
            sa, va = cpu_bench.radiosity_sample(solid, color, max(64, 8 * cores), cores)
            out["radiosity"]["cpu_baseline"] = {"value": va / sa / 1e6, "unit": "Mvoxel/s", "cores": cores, "kind": cpu_bench.kind(),
                                                "one_core_mvoxel_s": v1 / s1 / 1e6,
                                                "full_map_seconds_extrapolated": 512 * 512 * 64 / (va / sa),
                                                "sample": f"GLRadiosityRenderer::UpdateChunk over {va // 4096} of the 4096 chunks (even "
                                                          f"spread, all four z layers) on {cores} threads, {sa:.2f} s wall; 32 chunks on one"}
        except Exception as e:
            out["radiosity"]["cpu_baseline"] = {"value": None, "kind": "unavailable", "sample": repr(e)}
    return out


_JSON_FD = None


def claim_stdout():
    """stdout carries exactly one JSON line.  Libraries that write to file descriptor 1 on their own (NCCL prints its
    "NCCL version ..." banner there under torchrun) are pointed at stderr; the JSON line goes to the original stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    args = parse_args()
    claim_stdout()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
