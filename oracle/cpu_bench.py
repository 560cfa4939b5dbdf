"""Times the reference's own CPU code (oracle/_ref/liboref_sse2.so = unmodified yvt/openspades sources
compiled in place, see oracle/Makefile) or, when that library did not travel, the C restatement
(oracle/liboracle.so) on this machine's host cores.

TEST / BASELINE INFRASTRUCTURE ONLY: imported by bench.py's `cpu_baseline` leg and `--impl reference`
arm, never by the product (openspades_b200/).

The reference has no batch API: SWRenderer renders one view per StartScene/EndScene and caps its own
fork-join at 32 threads (Sources/Draw/SWUtils.h:53); GameMap::CastRay2 is one ray per call.  To give the
CPU "all the host threads it can use" the views / rays are split over P worker processes, each owning
a private GameMap + SWRenderer (r_swNumThreads = 1), started on a common wall-clock tick; throughput is
total units / (last worker's finish - tick).
"""
import ctypes
import json
import os
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

SCENE_DTYPE = np.dtype([("origin", "<f4", 3), ("axis", "<f4", (3, 3)), ("fovX", "<f4"), ("fovY", "<f4"),
                        ("zNear", "<f4"), ("zFar", "<f4")])


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def ref_lib_path(variant="sse2"):
    return os.path.join(HERE, "_ref", f"liboref_{variant}.so")


def kind():
    """'reference' when the compiled reference travelled to this machine, else 'port'."""
    return "reference" if os.path.exists(ref_lib_path()) else "port"


def _load_ref():
    vp, ci = ctypes.c_void_p, ctypes.c_int
    L = ctypes.CDLL(ref_lib_path())
    L.oref_map_from_arrays.restype = vp
    L.oref_map_from_arrays.argtypes = [vp, vp]
    L.oref_renderer_create.restype = vp
    L.oref_renderer_create.argtypes = [ci] * 4
    L.oref_renderer_set_map.argtypes = [vp, vp]
    L.oref_renderer_set_fog.argtypes = [vp] + [ctypes.c_float] * 4
    L.oref_renderer_render.argtypes = [vp, vp, vp, vp]
    L.oref_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
    return L


def _load_port():
    vp, ci = ctypes.c_void_p, ctypes.c_int
    path = os.path.join(HERE, "liboracle.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    L = ctypes.CDLL(path)
    L.orc_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
    L.orc_castray2.restype = None
    L.orc_render.argtypes = [vp, vp, vp, ci, ci, vp, ctypes.c_float, ci, ci, ci, vp, vp]
    return L


def _worker(spec_path, idx):
    # the reference logs (SPLog) to stdout: keep a private copy of the pipe for the protocol and send
    # everything else written to fd 1 to /dev/null
    proto = os.fdopen(os.dup(1), "w")
    os.dup2(os.open(os.devnull, os.O_WRONLY), 1)
    spec = json.load(open(spec_path))
    d = np.load(spec["data"])
    solid, color = np.ascontiguousarray(d["solid"]), np.ascontiguousarray(d["color"])
    use_ref = spec["kind"] == "reference"
    L = _load_ref() if use_ref else _load_port()
    lo, hi = spec["ranges"][idx]
    m = L.oref_map_from_arrays(solid.ctypes.data, color.ctypes.data) if use_ref else None
    if spec["what"] == "views":
        w, h = spec["w"], spec["h"]
        scenes = np.ascontiguousarray(d["scenes"][lo:hi])
        fog = spec["fog"]
        fb = np.zeros((h, w), np.uint32)
        dp = np.zeros((h, w), np.float32)
        if use_ref:
            r = L.oref_renderer_create(w, h, -1, spec["sw_threads"])
            L.oref_renderer_set_map(r, m)
            L.oref_renderer_set_fog(r, fog[0], fog[1], fog[2], spec["fog_dist"])

            def run(i):
                L.oref_renderer_render(r, scenes[i:i + 1].ctypes.data, fb.ctypes.data, dp.ctypes.data)
        else:
            fc = np.asarray(fog, np.float32)

            def run(i):
                L.orc_render(solid.ctypes.data, color.ctypes.data, scenes[i:i + 1].ctypes.data, w, h, fc.ctypes.data,
                             spec["fog_dist"], 2, spec["sw_threads"], 1, fb.ctypes.data, dp.ctypes.data)
        if len(scenes):
            run(0)  # warm-up (page in, RLE build done in set_map)
        n = len(scenes)
    else:  # rays
        v0 = np.ascontiguousarray(d["v0"][lo:hi])
        dr = np.ascontiguousarray(d["dir"][lo:hi])
        out = np.zeros((hi - lo) * 40, np.uint8)
        n = hi - lo

        def run_all():
            if use_ref:
                L.oref_castray2(m, v0.ctypes.data, dr.ctypes.data, spec["max_steps"], out.ctypes.data, n, 1)
            else:
                L.orc_castray2(solid.ctypes.data, v0.ctypes.data, dr.ctypes.data, spec["max_steps"], out.ctypes.data, n, 1)
    print("ready", file=proto, flush=True)
    for line in sys.stdin:  # one "go <t0>" per step
        parts = line.split()
        if not parts or parts[0] != "go":
            break
        t0 = float(parts[1])
        while time.time() < t0:
            time.sleep(0.0005)
        t_start = time.time()
        if spec["what"] == "views":
            for i in range(n):
                run(i)
        else:
            run_all()
        t_end = time.time()
        print(json.dumps({"idx": idx, "n": n, "start": t_start, "end": t_end}), file=proto, flush=True)


def single_view_ms(solid, color, scene, w, h, fog, fog_dist, threads=4, reps=10):
    """One view, the way the game renders it: a single SWRenderer with r_swNumThreads = threads (default 4,
    SWRenderer.cpp:41); best of `reps` wall times around StartScene .. EndScene .. FrameDone, in ms."""
    solid, color = np.ascontiguousarray(solid), np.ascontiguousarray(color)
    sc = np.ascontiguousarray(np.atleast_1d(scene))
    fb = np.zeros((h, w), np.uint32)
    dp = np.zeros((h, w), np.float32)
    if kind() == "reference":
        L = _load_ref()
        m = L.oref_map_from_arrays(solid.ctypes.data, color.ctypes.data)
        r = L.oref_renderer_create(w, h, -1, threads)
        L.oref_renderer_set_map(r, m)
        L.oref_renderer_set_fog(r, fog[0], fog[1], fog[2], fog_dist)

        def run():
            L.oref_renderer_render(r, sc.ctypes.data, fb.ctypes.data, dp.ctypes.data)
    else:
        L = _load_port()
        fc = np.asarray(fog, np.float32)

        def run():
            L.orc_render(solid.ctypes.data, color.ctypes.data, sc.ctypes.data, w, h, fc.ctypes.data, fog_dist, 2, threads, 1,
                         fb.ctypes.data, dp.ctypes.data)
    run()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        run()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3


def radiosity_sample(solid, color, n_chunks=64, threads=1):
    """GLRadiosityRenderer::UpdateChunk (Evaluate x 4096 + EncodeValue x 4, GLRadiosityRenderer.cpp:517-543) over a
    deterministic spread of `n_chunks` of the 4096 chunks of the map, `threads` chunks at a time (the function is
    independent per chunk).  Returns (seconds, voxels)."""
    from concurrent.futures import ThreadPoolExecutor
    solid, color = np.ascontiguousarray(solid), np.ascontiguousarray(color)
    # chunk index = (cx + cy * 32) * 4 + cz: an even spread over the map that also cycles through the four z layers
    ids = [((i * 4096 // n_chunks) & ~3) | (i & 3) for i in range(n_chunks)]
    todo = [((ci >> 2) & 31, ci >> 7, ci & 3) for ci in ids]
    vp, ci_ = ctypes.c_void_p, ctypes.c_int
    if kind() == "reference":
        L = _load_ref()
        L.oref_rad_create.restype = vp
        L.oref_rad_create.argtypes = [vp, ci_]
        L.oref_rad_shadow_update.argtypes = [vp]
        L.oref_rad_update_chunk.argtypes = [vp, ci_, ci_, ci_]
        m = L.oref_map_from_arrays(solid.ctypes.data, color.ctypes.data)
        hnd = L.oref_rad_create(m, 1)
        L.oref_rad_shadow_update(hnd)

        def one(c):
            L.oref_rad_update_chunk(hnd, c[0], c[1], c[2])
    else:
        L = _load_port()
        L.orc_shadow_bitmap.argtypes = [vp, vp, vp]
        L.orc_shadow_bitmap.restype = None
        L.orc_radiosity_box.argtypes = [vp] + [ci_] * 7 + [vp, vp, ci_]
        L.orc_radiosity_box.restype = None
        bitmap = np.zeros((512, 512), np.uint32)
        L.orc_shadow_bitmap(solid.ctypes.data, color.ctypes.data, bitmap.ctypes.data)
        outs = {}

        def one(c):
            enc = outs.setdefault(c, np.zeros(4096 * 4, np.uint32))
            L.orc_radiosity_box(bitmap.ctypes.data, c[0] * 16, c[1] * 16, c[2] * 16, c[0] * 16 + 15, c[1] * 16 + 15,
                                c[2] * 16 + 15, 1, None, enc.ctypes.data, 1)
    one(todo[0])
    t0 = time.perf_counter()
    if threads <= 1:
        for c in todo:
            one(c)
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(one, todo))
    return time.perf_counter() - t0, len(todo) * 4096


class Pool:
    """P persistent worker processes, each with a private copy of the reference objects and a fixed share
    of the units; step() runs every share once from a common wall-clock tick."""

    def __init__(self, what, solid, color, units, workers=None, tmpdir="/tmp", **kw):
        workers = workers or usable_cores()
        n = len(units) if what == "views" else len(units[0])
        self.workers = workers = max(1, min(workers, n))
        self.kind = kind()
        os.makedirs(tmpdir, exist_ok=True)
        self.data = os.path.join(tmpdir, f"osb_cpu_bench_{os.getpid()}.npz")
        if what == "views":
            np.savez(self.data, solid=solid, color=color, scenes=units)
        else:
            np.savez(self.data, solid=solid, color=color, v0=units[0], dir=units[1])
        bounds = [n * i // workers for i in range(workers + 1)]
        spec = dict(kw, what=what, kind=self.kind, data=self.data,
                    ranges=[[bounds[i], bounds[i + 1]] for i in range(workers)])
        self.spec_path = os.path.join(tmpdir, f"osb_cpu_bench_{os.getpid()}.json")
        json.dump(spec, open(self.spec_path, "w"))
        self.procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", self.spec_path, str(i)],
                                       stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, bufsize=1)
                      for i in range(workers)]
        for p in self.procs:
            line = p.stdout.readline()
            if line.strip() != "ready":
                raise RuntimeError(f"cpu_bench worker failed to start: {line!r}")

    def step(self):
        """Runs every worker's share once; returns dict(seconds, units)."""
        t0 = time.time() + 0.05 + 0.001 * self.workers
        for p in self.procs:
            p.stdin.write(f"go {t0!r}\n")
            p.stdin.flush()
        res = []
        for p in self.procs:
            line = p.stdout.readline()
            if not line:
                raise RuntimeError("cpu_bench worker died")
            res.append(json.loads(line))
        return {"seconds": max(r["end"] for r in res) - t0, "units": sum(r["n"] for r in res),
                "late_start_s": max(r["start"] for r in res) - t0}

    def close(self):
        for p in self.procs:
            try:
                p.stdin.close()
            except OSError:
                pass
        for p in self.procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
        for f in (self.data, self.spec_path):
            if os.path.exists(f):
                os.remove(f)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


if __name__ == "__main__":
    if len(sys.argv) >= 4 and sys.argv[1] == "--worker":
        _worker(sys.argv[2], int(sys.argv[3]))
