#!/usr/bin/env python
"""The box's device->host ceiling: N concurrent 1 GiB copies from N GPUs into pinned host memory, nothing else running.

bench.py's e2e at N GPUs returns 1 GiB of BGRX frames per GPU per step; what it can reach is bounded by this figure, not
by anything on the GPU side.  Launch like the bench:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/d2h_ceiling.py
Rank 0 prints one JSON line (and writes gpurun_out/r02_d2h_ceiling_nN.json).  With --affinity each rank first binds itself
to the CPUs nvidia-smi reports as nearest to its GPU (the pinned allocation then lands on that NUMA node by first touch).
"""
import argparse
import json
import os
import subprocess
import time


def gpu_cpu_affinity(index):
    """CPU list nearest GPU `index` from `nvidia-smi topo -C -i` (e.g. '0-55,112-167'); None when the tool says nothing."""
    try:
        out = subprocess.check_output(["nvidia-smi", "topo", "-C", "-i", str(index)], text=True, stderr=subprocess.DEVNULL)
    except Exception:
        return None
    for line in out.splitlines():
        if ":" in line:
            spec = line.split(":")[-1].strip()
            cpus = set()
            try:
                for part in spec.split(","):
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
            except ValueError:
                continue
            if cpus:
                return cpus
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=1.0)
    ap.add_argument("--reps", type=int, default=8)
    ap.add_argument("--affinity", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    bound = None
    if args.affinity:
        cpus = gpu_cpu_affinity(local)
        if cpus:
            allowed = cpus & os.sched_getaffinity(0)
            if allowed:
                os.sched_setaffinity(0, allowed)
                bound = len(allowed)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = int(args.gib * (1 << 30))
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    d.fill_(rank + 1)
    h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h.fill_(0)  # first touch on this rank's (possibly bound) CPUs
    for _ in range(2):
        h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()

    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.reps):
        h.copy_(d, non_blocking=True)
    e1.record()
    e1.synchronize()
    barrier()
    wall = time.perf_counter() - t0
    mine = n * args.reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    ok = bool(int(h[12345]) == rank + 1)
    per_rank = [mine]
    if world > 1:
        t = torch.tensor([mine], device="cuda")
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        per_rank = [float(x) for x in g]
    if rank == 0:
        topo = ""
        try:
            topo = subprocess.check_output(["nvidia-smi", "topo", "-m"], text=True, stderr=subprocess.DEVNULL)
        except Exception:
            pass
        line = {"what": f"{world} concurrent D2H copies of {args.gib} GiB each into pinned host memory, {args.reps} back to back",
                "n_gpus": world, "per_gpu_gb_s": per_rank, "sum_gb_s_device_timed": sum(per_rank),
                "aggregate_gb_s_wall": world * n * args.reps / wall / 1e9, "cpu_affinity": "nearest to the GPU" if args.affinity else "none",
                "cpus_bound_rank0": bound, "copied_ok": ok, "host_cpus": os.cpu_count(), "topo": topo.splitlines()[:14]}
        print(json.dumps(line))
        os.makedirs("gpurun_out", exist_ok=True)
        tag = f"n{world}" + ("_affinity" if args.affinity else "")
        with open(os.path.join("gpurun_out", f"r02_d2h_ceiling_{tag}.json"), "w") as f:
            f.write(json.dumps(line) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
