"""Renders the bench's in-game 1080p frame (terrain + 64 model instances + 8 lights + fog) a few times; the command
tools/gpu_game_frame.sh profiles.  Prints per-kernel CUDA-event times of a serialised pass."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap
n_inst = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_lights = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
m = VoxelMap.load_file(workloads.TITLE_VXL)
w, h = 1920, 1080
with osb.Context(0, osb.NUM_SSE2, sw_threads=4) as ctx:
    ctx.map_upload(m.solid.copy(), m.color.copy())
    sc = workloads.title_scene(w, h, 68.0)
    handles = [ctx.model_create(*workloads.blob_model(11 + k)) for k in range(3)]
    inst, lights = workloads.game_frame(sc, handles, n_inst, n_lights)
    ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, to_host=False)
    ctx.timing_enable(True)
    ctx.timing_read()
    for _ in range(reps):
        ctx.render_scene(sc, w, h, (0.1, 0.1, 0.1), 128.0, inst, lights, to_host=False)
    for k, (n, ms) in ctx.timing_read().items():
        if n:
            print(f"{k:18s} {n:4d} launches  {ms / reps:8.4f} ms/frame")
    for hd in handles:
        ctx.model_destroy(hd)
