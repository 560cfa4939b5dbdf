"""ctypes binding of the osb200 C ABI (include/osb200.h) -- used by tests/, bench.py and smoke().

The product is the CUDA library; this file only marshals numpy / torch buffers into the plain
pointers the C ABI takes.  Loading fails loudly when the library has not been built, and every
compute call fails loudly (OsbError) when no CUDA device is present: there is no CPU fallback.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("OSB200_LIB") or os.path.join(_HERE, "lib", "libosb200.so")  # override: kernel A/B experiments

NUM_IEEE, NUM_SSE, NUM_SSE2 = 0, 1, 2
# osb200_model_instance (pointer + 16 + 3 floats, 4 bytes of tail padding) and osb200_light
MODEL_INSTANCE_DTYPE = np.dtype({"names": ["model", "matrix", "customColor"], "formats": ["<u8", ("<f4", 16), ("<f4", 3)],
                                 "offsets": [0, 8, 72], "itemsize": 88})
LIGHT_DTYPE = np.dtype([("origin", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 3)])


class OsbError(RuntimeError):
    pass


class Opts(ctypes.Structure):
    _fields_ = [
        ("device", ctypes.c_int),
        ("numerics", ctypes.c_int),
        ("sw_threads", ctypes.c_int),
        ("undersampling", ctypes.c_int),
        ("radiosity_mode", ctypes.c_int),
        ("rcp_table", ctypes.c_void_p),
        ("rcp_bits", ctypes.c_int),
        ("rsqrt_table", ctypes.c_void_p),
        ("rsqrt_bits", ctypes.c_int),
    ]


class KernelTime(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("launches", ctypes.c_uint64), ("total_ms", ctypes.c_double)]


# numpy dtypes matching the C structs
SCENE_DTYPE = np.dtype([("origin", "<f4", 3), ("axis", "<f4", (3, 3)), ("fovX", "<f4"), ("fovY", "<f4"),
                        ("zNear", "<f4"), ("zFar", "<f4")])
RAYHIT_DTYPE = np.dtype([("hit", "u1"), ("startSolid", "u1"), ("pad", "u1", 2), ("hitPos", "<f4", 3),
                         ("hitBlock", "<i4", 3), ("normal", "<i4", 3)])
assert SCENE_DTYPE.itemsize == 64 and RAYHIT_DTYPE.itemsize == 40

_lib = None


def load_library():
    """dlopen libosb200.so; raises OsbError (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OsbError(f"{LIB_PATH} not built: run `python -m openspades_b200.build` (needs nvcc); "
                       "there is no CPU fallback")
    lib = ctypes.CDLL(LIB_PATH)
    vp, ci, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    lib.osb200_last_error.restype = ctypes.c_char_p
    lib.osb200_create.argtypes = [ctypes.POINTER(Opts), ctypes.POINTER(vp)]
    lib.osb200_destroy.argtypes = [vp]
    lib.osb200_destroy.restype = None
    lib.osb200_sync.argtypes = [vp]
    lib.osb200_stream.argtypes = [vp]
    lib.osb200_stream.restype = vp
    lib.osb200_launch_count.argtypes = [vp]
    lib.osb200_launch_count.restype = ctypes.c_uint64
    lib.osb200_timing_enable.argtypes = [vp, ci]
    lib.osb200_timing_read.argtypes = [vp, ctypes.POINTER(KernelTime), ci]
    lib.osb200_map_upload.argtypes = [vp, vp, vp]
    lib.osb200_map_clear.argtypes = [vp]
    lib.osb200_map_load_vxl.argtypes = [vp, vp, sz, vp, vp]
    lib.osb200_map_save_vxl.argtypes = [vp, vp, sz]
    lib.osb200_map_save_vxl.restype = ctypes.c_longlong
    lib.osb200_links_rebuild.argtypes = [vp]
    lib.osb200_links_blocks_added.argtypes = [vp]
    lib.osb200_links_remove_blocks.argtypes = [vp, vp, ci, vp, ci]
    lib.osb200_links_read.argtypes = [vp, vp]
    lib.osb200_links_last_floating.argtypes = [vp, vp, ci]
    lib.osb200_links_last_rounds.argtypes = [vp]
    lib.osb200_clusterize_blocks.argtypes = [vp, vp, ci, vp]
    lib.osb200_minimap.argtypes = [vp, vp, ctypes.POINTER(vp)]
    lib.osb200_map_mark_dirty.argtypes = [vp, ci, ci, ci]
    lib.osb200_map_flush.argtypes = [vp, vp, vp]
    lib.osb200_map_read_solid.argtypes = [vp, vp]
    lib.osb200_castray2.argtypes = [vp, vp, vp, ci, vp, sz]
    lib.osb200_castray2_device.argtypes = [vp, vp, vp, ci, vp, sz]
    lib.osb200_castray2_static.argtypes = [vp, vp, vp, ci, vp, sz]
    fp = ctypes.c_float
    lib.osb200_render.argtypes = [vp, vp, ci, ci, ci, vp, fp, vp, vp]
    lib.osb200_render_device.argtypes = [vp, vp, ci, ci, ci, vp, fp, vp, vp]
    lib.osb200_read_depth.argtypes = [vp, vp, sz]
    lib.osb200_shadow_bitmap.argtypes = [vp, vp]
    lib.osb200_radiosity_update.argtypes = [vp, ci, vp]
    lib.osb200_radiosity_chunk_info.argtypes = [vp, ci, ci, ci, vp]
    lib.osb200_radiosity_read_chunk.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp]
    lib.osb200_radiosity_evaluate_box.argtypes = [vp] + [ci] * 6 + [vp]
    lib.osb200_radiosity_volumes.argtypes = [vp] + [ctypes.POINTER(vp)] * 5
    lib.osb200_radiosity_count_taps.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]
    lib.osb200_castray.argtypes = [vp, vp, vp, fp, vp, sz]
    lib.osb200_model_create.argtypes = [vp, ci, ci, ci, vp, vp, vp, ctypes.POINTER(vp)]
    lib.osb200_model_destroy.argtypes = [vp]
    lib.osb200_model_destroy.restype = None
    lib.osb200_model_render_data.argtypes = [vp, vp, ci, vp, ci]
    lib.osb200_render_scene.argtypes = [vp, vp, ci, ci, vp, fp, vp, ci, vp, ci, vp, vp]
    lib.osb200_castray_device.argtypes = [vp, vp, vp, fp, vp, sz]
    lib.osb200_ambient_set_rays.argtypes = [vp, vp]
    lib.osb200_ambient_get_rays.argtypes = [vp, vp]
    lib.osb200_ambient_update.argtypes = [vp, ci, vp]
    lib.osb200_ambient_chunk_info.argtypes = [vp, ci, ci, ci, vp]
    lib.osb200_ambient_read_chunk.argtypes = [vp, ci, ci, ci, vp]
    lib.osb200_ambient_evaluate_box.argtypes = [vp] + [ci] * 6 + [vp]
    lib.osb200_ambient_volume.argtypes = [vp, ctypes.POINTER(vp)]
    lib.osb200_group_create.argtypes = [ctypes.POINTER(Opts), ctypes.POINTER(ci), ci, ctypes.POINTER(vp)]
    lib.osb200_group_destroy.argtypes = [vp]
    lib.osb200_group_destroy.restype = None
    lib.osb200_group_size.argtypes = [vp]
    lib.osb200_group_ctx.argtypes = [vp, ci]
    lib.osb200_group_ctx.restype = vp
    lib.osb200_group_map_upload.argtypes = [vp, vp, vp]
    lib.osb200_group_map_mark_dirty.argtypes = [vp, ci, ci, ci]
    lib.osb200_group_map_flush.argtypes = [vp, vp, vp]
    lib.osb200_group_render.argtypes = [vp, vp, ci, ci, ci, vp, fp, vp, vp]
    lib.osb200_group_render_gather.argtypes = [vp, vp, ci, ci, ci, vp, fp, ci, vp, vp]
    lib.osb200_group_castray2.argtypes = [vp, vp, vp, ci, vp, sz]
    _lib = lib
    return lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):  # torch tensor (device or pinned host memory)
        return a.data_ptr()
    return a


def calibrate_sse(max_bits=16):
    """Host RCPSS/RSQRTSS tables (osb200_calibrate_sse). Pure host code; works without a GPU."""
    lib = load_library()
    lib.osb200_calibrate_sse.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.c_void_p,
                                         ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    rcp = np.zeros(1 << max_bits, np.uint32)
    rsq = np.zeros(2 << max_bits, np.uint32)
    rb, sb = ctypes.c_int(0), ctypes.c_int(0)
    rc = lib.osb200_calibrate_sse(rcp.ctypes.data, ctypes.byref(rb), rsq.ctypes.data, ctypes.byref(sb), max_bits)
    if rc:
        raise OsbError(lib.osb200_last_error().decode())
    return rcp[: 1 << rb.value].copy(), rb.value, rsq[: 2 << sb.value].copy(), sb.value


class Context:
    """Owns one osb200_ctx (one GPU).  Mirrors the C ABI one to one."""

    def __init__(self, device=0, numerics=NUM_SSE2, sw_threads=4, undersampling=1, radiosity_mode=1):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        o = Opts(device=device, numerics=numerics, sw_threads=sw_threads, undersampling=undersampling,
                 radiosity_mode=radiosity_mode, rcp_table=None, rcp_bits=0, rsqrt_table=None, rsqrt_bits=0)
        rc = self._lib.osb200_create(ctypes.byref(o), ctypes.byref(self._h))
        if rc:
            raise OsbError(f"osb200_create failed ({rc}): {self._lib.osb200_last_error().decode()}")
        self.device = device
        self.numerics = numerics

    def _check(self, rc):
        if rc < 0:
            raise OsbError(f"osb200 error {rc}: {self._lib.osb200_last_error().decode()}")
        return rc

    def close(self):
        if getattr(self, "_h", None):
            self._lib.osb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def stream(self):
        return self._lib.osb200_stream(self._h)

    @property
    def launch_count(self):
        return int(self._lib.osb200_launch_count(self._h))

    def sync(self):
        self._check(self._lib.osb200_sync(self._h))

    def timing_enable(self, on=True):
        self._check(self._lib.osb200_timing_enable(self._h, int(bool(on))))

    def timing_read(self):
        """{kernel name: (launches, total_ms)} since the last read (CUDA events on the context stream)."""
        arr = (KernelTime * 32)()
        n = self._check(self._lib.osb200_timing_read(self._h, arr, 32))
        return {arr[i].name.decode(): (int(arr[i].launches), float(arr[i].total_ms)) for i in range(n)}

    # ---- map ----
    def map_upload(self, solid, color):
        solid = np.ascontiguousarray(solid, np.uint64)
        color = np.ascontiguousarray(color, np.uint32)
        assert solid.shape == (512, 512) and color.shape == (512, 512, 64)
        self._check(self._lib.osb200_map_upload(self._h, solid.ctypes.data, color.ctypes.data))

    def map_load_vxl(self, data: bytes, want_arrays=True):
        """GameMap::Load on the device: raw VXL bytes -> resident map; returns (solid, color) host arrays or None."""
        buf = np.frombuffer(data, np.uint8)
        solid = np.empty((512, 512), np.uint64) if want_arrays else None
        color = np.empty((512, 512, 64), np.uint32) if want_arrays else None
        self._check(self._lib.osb200_map_load_vxl(self._h, buf.ctypes.data, len(buf), _ptr(solid), _ptr(color)))
        return (solid, color) if want_arrays else None

    def map_save_vxl(self) -> bytes:
        """GameMap::Save from the resident map."""
        n = self._lib.osb200_map_save_vxl(self._h, None, 0)
        if n < 0:
            raise OsbError(self._lib.osb200_last_error().decode())
        out = np.empty(n, np.uint8)
        m = self._lib.osb200_map_save_vxl(self._h, out.ctypes.data, n)
        if m != n:
            raise OsbError(self._lib.osb200_last_error().decode() if m < 0 else "osb200_map_save_vxl: length changed")
        return out.tobytes()

    # ---- GameMapWrapper (floating blocks) + minimap ----
    def links_rebuild(self):
        return self._check(self._lib.osb200_links_rebuild(self._h))

    def links_blocks_added(self):
        return self._check(self._lib.osb200_links_blocks_added(self._h))

    def links_remove_blocks(self, cells, cap=4096):
        """cells: (n, 3) int16 already cleared + flushed; returns the (m, 3) int16 floating blocks sorted by (x, y, z)."""
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros((cap, 3), np.int16)
        n = self._check(self._lib.osb200_links_remove_blocks(self._h, cells.ctypes.data, len(cells), out.ctypes.data, cap))
        if n > cap:
            out = np.zeros((n, 3), np.int16)
            self._check(self._lib.osb200_links_last_floating(self._h, out.ctypes.data, n))
        return out[:n].copy()

    def links_read(self):
        out = np.zeros((512, 512), np.uint64)
        self._check(self._lib.osb200_links_read(self._h, out.ctypes.data))
        return out

    def clusterize_blocks(self, cells):
        """World::ClusterizeBlocks (Client/World.cpp:154-226): cells int16 [n, 3] -> (count, int32 [n] component number
        of every cell, components numbered by their smallest cell)."""
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros(len(cells), np.int32)
        k = self._check(self._lib.osb200_clusterize_blocks(self._h, cells.ctypes.data, len(cells), out.ctypes.data))
        return k, out

    def links_last_rounds(self):
        return self._lib.osb200_links_last_rounds(self._h)

    def minimap(self):
        out = np.zeros((512, 512), np.uint32)
        self._check(self._lib.osb200_minimap(self._h, out.ctypes.data, None))
        return out

    def map_clear(self):
        self._check(self._lib.osb200_map_clear(self._h))

    def map_mark_dirty(self, x, y, z):
        self._check(self._lib.osb200_map_mark_dirty(self._h, int(x), int(y), int(z)))

    def map_flush(self, solid, color):
        assert solid.dtype == np.uint64 and color.dtype == np.uint32
        assert solid.flags.c_contiguous and color.flags.c_contiguous
        return self._check(self._lib.osb200_map_flush(self._h, solid.ctypes.data, color.ctypes.data))

    def map_read_solid(self):
        out = np.zeros((512, 512), np.uint64)
        self._check(self._lib.osb200_map_read_solid(self._h, out.ctypes.data))
        return out

    # ---- GameMap::CastRay2 ----
    def castray2(self, v0, dirs, max_steps, out=None):
        """Host arrays in, host structured array (RAYHIT_DTYPE) out.  v0 / dirs / out may be pinned-memory buffers (e.g.
        torch pin_memory tensors viewed as numpy): the chunk pipeline behind the call then runs at PCIe speed."""
        if hasattr(v0, "data_ptr"):
            n = v0.shape[0]
            pv, pd = v0.data_ptr(), dirs.data_ptr()
        else:
            v0 = np.ascontiguousarray(v0, np.float32)
            dirs = np.ascontiguousarray(dirs, np.float32)
            n = v0.shape[0]
            assert v0.shape == (n, 3) and dirs.shape == (n, 3)
            pv, pd = v0.ctypes.data, dirs.ctypes.data
        if out is None:
            out = np.zeros(n, RAYHIT_DTYPE)
        self._check(self._lib.osb200_castray2(self._h, pv, pd, int(max_steps), _ptr(out), n))
        return out

    CASTRAY2_STATIC_MAX_RAYS = 65536  # OSB200_CASTRAY2_STATIC_MAX_RAYS (include/osb200.h)

    def castray2_static(self, v0, dirs, max_steps, out=None):
        """osb200_castray2_static: same arguments and results as castray2, no allocation inside the call (static device
        scratch, at most CASTRAY2_STATIC_MAX_RAYS rays per call; calls that overlap on one GPU race)."""
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = v0.shape[0]
        assert v0.shape == (n, 3) and dirs.shape == (n, 3)
        if out is None:
            out = np.zeros(n, RAYHIT_DTYPE)
        self._check(self._lib.osb200_castray2_static(self._h, v0.ctypes.data, dirs.ctypes.data, int(max_steps), _ptr(out), n))
        return out

    def castray2_device(self, d_v0, d_dir, max_steps, d_out, n):
        """Device pointers (ints or torch CUDA tensors); enqueue only."""
        self._check(self._lib.osb200_castray2_device(self._h, _ptr(d_v0), _ptr(d_dir), int(max_steps),
                                                     _ptr(d_out), int(n)))

    # ---- terrain pass + fog ----
    def render(self, scenes, w, h, fog_color, fog_dist, want_depth=True, out_bgrx=None, out_depth=None):
        """Host entry point: scenes = SCENE_DTYPE array (n,), returns (bgrx[n,h,w] uint32, depth[n,h,w] f32)."""
        scenes = np.ascontiguousarray(np.atleast_1d(scenes))
        assert scenes.dtype == SCENE_DTYPE
        n = len(scenes)
        fb = np.empty((n, h, w), np.uint32) if out_bgrx is None else out_bgrx
        dp = (np.empty((n, h, w), np.float32) if out_depth is None else out_depth) if want_depth else None
        fc = np.asarray(fog_color, np.float32)
        self._check(self._lib.osb200_render(self._h, scenes.ctypes.data, n, w, h, fc.ctypes.data, float(fog_dist),
                                            _ptr(fb), _ptr(dp)))
        return fb, dp

    def render_device(self, scenes, w, h, fog_color, fog_dist, d_bgrx, d_depth):
        """Device outputs (torch CUDA tensors or raw pointers); enqueue only."""
        scenes = np.ascontiguousarray(np.atleast_1d(scenes))
        assert scenes.dtype == SCENE_DTYPE
        fc = np.asarray(fog_color, np.float32)
        self._check(self._lib.osb200_render_device(self._h, scenes.ctypes.data, len(scenes), w, h, fc.ctypes.data,
                                                   float(fog_dist), _ptr(d_bgrx), _ptr(d_depth)))

    # ---- shadow bitmap + radiosity ----
    def shadow_bitmap(self):
        out = np.zeros((512, 512), np.uint32)  # [y][x]
        self._check(self._lib.osb200_shadow_bitmap(self._h, out.ctypes.data))
        return out

    def radiosity_update(self, max_chunks=-1, eye=None):
        e = None if eye is None else np.ascontiguousarray(eye, np.float32)
        return self._check(self._lib.osb200_radiosity_update(self._h, int(max_chunks), None if e is None else e.ctypes.data))

    def radiosity_count_taps(self):
        """(executed, total) taps of a full-map Evaluate: measurement hook for the FP32 roofline of C5"""
        a, t = ctypes.c_ulonglong(0), ctypes.c_ulonglong(0)
        self._check(self._lib.osb200_radiosity_count_taps(self._h, ctypes.byref(a), ctypes.byref(t)))
        return int(a.value), int(t.value)

    def radiosity_chunk_info(self, cx, cy, cz):
        out = np.zeros(7, np.int32)
        self._check(self._lib.osb200_radiosity_chunk_info(self._h, cx, cy, cz, out.ctypes.data))
        return out

    def radiosity_read_chunk(self, cx, cy, cz):
        """(16,16,16,4) uint32, [z][y][x][flat,x,y,z]"""
        a = [np.zeros((16, 16, 16), np.uint32) for _ in range(4)]
        self._check(self._lib.osb200_radiosity_read_chunk(self._h, cx, cy, cz, *[v.ctypes.data for v in a]))
        return np.stack(a, axis=-1)

    def radiosity_evaluate_box(self, lo, hi):
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        out = np.zeros((nz, ny, nx, 12), np.float32)
        self._check(self._lib.osb200_radiosity_evaluate_box(self._h, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                                                            out.ctypes.data))
        return out

    # ---- voxel models + dynamic lights: the full EndScene ----
    def model_create(self, w, h, d, origin, solid, colors):
        """SWModel from VoxelModel arrays: solid uint64[h][w] (index x + y*w), colors uint32[h][w][d]. Returns a handle."""
        o = np.asarray(origin, np.float32)
        solid = np.ascontiguousarray(solid, np.uint64)
        colors = np.ascontiguousarray(colors, np.uint32)
        assert solid.size == w * h and colors.size == w * h * d
        m = ctypes.c_void_p()
        self._check(self._lib.osb200_model_create(self._h, w, h, d, o.ctypes.data, solid.ctypes.data, colors.ctypes.data,
                                                  ctypes.byref(m)))
        return m.value

    def model_destroy(self, m):
        self._lib.osb200_model_destroy(m)

    def model_render_data(self, m, ncols):
        n = self._lib.osb200_model_render_data(m, None, 0, None, 0)
        data = np.zeros(n, np.uint32)
        addr = np.zeros(ncols, np.uint32)
        self._lib.osb200_model_render_data(m, data.ctypes.data, n, addr.ctypes.data, ncols)
        return data, addr

    def render_scene(self, scene, w, h, fog_color, fog_dist, instances=None, lights=None, want_depth=True,
                     out_bgrx=None, out_depth=None, to_host=True):
        """One view: terrain + models (MODEL_INSTANCE_DTYPE[n]) + point lights (LIGHT_DTYPE[m]) + fog.
        out_bgrx / out_depth: host tensors or arrays to fill (pinned for speed); to_host=False leaves the frame in
        the context's device buffers (timing of the device leg only)."""
        scene = np.ascontiguousarray(np.atleast_1d(scene))
        assert scene.dtype == SCENE_DTYPE and len(scene) == 1
        inst = np.zeros(0, MODEL_INSTANCE_DTYPE) if instances is None else np.ascontiguousarray(instances)
        lg = np.zeros(0, LIGHT_DTYPE) if lights is None else np.ascontiguousarray(lights)
        assert inst.dtype == MODEL_INSTANCE_DTYPE and lg.dtype == LIGHT_DTYPE
        fb = dp = None
        if to_host:
            fb = np.empty((h, w), np.uint32) if out_bgrx is None else out_bgrx
            dp = out_depth if out_depth is not None else (np.empty((h, w), np.float32) if want_depth else None)
        fc = np.asarray(fog_color, np.float32)
        self._check(self._lib.osb200_render_scene(self._h, scene.ctypes.data, w, h, fc.ctypes.data, float(fog_dist),
                                                  inst.ctypes.data if len(inst) else None, len(inst),
                                                  lg.ctypes.data if len(lg) else None, len(lg),
                                                  None if fb is None else _ptr(fb), None if dp is None else _ptr(dp)))
        return fb, dp

    # ---- GameMap::CastRay + ambient occlusion ----
    def castray(self, v0, dirs, length):
        """GameMap::CastRay batch: host arrays in, (n,4) int32 out: hit, hitBlock.xyz."""
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = v0.shape[0]
        assert v0.shape == (n, 3) and dirs.shape == (n, 3)
        out = np.zeros((n, 4), np.int32)
        self._check(self._lib.osb200_castray(self._h, v0.ctypes.data, dirs.ctypes.data, float(length), out.ctypes.data, n))
        return out

    def castray_device(self, d_v0, d_dir, length, d_out, n):
        self._check(self._lib.osb200_castray_device(self._h, _ptr(d_v0), _ptr(d_dir), float(length), _ptr(d_out), int(n)))

    def ambient_set_rays(self, rays):
        rays = np.ascontiguousarray(rays, np.float32)
        assert rays.size == 48
        self._check(self._lib.osb200_ambient_set_rays(self._h, rays.ctypes.data))

    def ambient_rays(self):
        out = np.zeros((16, 3), np.float32)
        self._check(self._lib.osb200_ambient_get_rays(self._h, out.ctypes.data))
        return out

    def ambient_update(self, max_chunks=-1, eye=None):
        e = None if eye is None else np.ascontiguousarray(eye, np.float32)
        return self._check(self._lib.osb200_ambient_update(self._h, int(max_chunks), None if e is None else e.ctypes.data))

    def ambient_chunk_info(self, cx, cy, cz):
        out = np.zeros(7, np.int32)
        self._check(self._lib.osb200_ambient_chunk_info(self._h, cx, cy, cz, out.ctypes.data))
        return out

    def ambient_read_chunk(self, cx, cy, cz):
        """(16,16,16,2) float32, [z][y][x][ao, air]"""
        out = np.zeros((16, 16, 16, 2), np.float32)
        self._check(self._lib.osb200_ambient_read_chunk(self._h, cx, cy, cz, out.ctypes.data))
        return out

    def ambient_evaluate_box(self, lo, hi):
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        out = np.zeros((nz, ny, nx), np.float32)
        self._check(self._lib.osb200_ambient_evaluate_box(self._h, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                                                          out.ctypes.data))
        return out

    def ambient_volume(self):
        """Device pointer (int) of the float2[65][512][512] volume."""
        p = ctypes.c_void_p()
        self._check(self._lib.osb200_ambient_volume(self._h, ctypes.byref(p)))
        return p.value

    def radiosity_volumes(self):
        """Device pointers (ints): flat, x, y, z volumes ([z][y][x] uint32) and the shadow bitmap."""
        p = [ctypes.c_void_p() for _ in range(5)]
        self._check(self._lib.osb200_radiosity_volumes(self._h, *[ctypes.byref(v) for v in p]))
        return [v.value for v in p]


class Group:
    """osb200_group: one context per listed GPU of this process, map replicated (PCIe once + NCCL broadcast), views and
    rays sharded by contiguous range.  Mirrors the osb200_group_* C entry points one to one."""

    def __init__(self, devices, numerics=NUM_SSE2, sw_threads=4, undersampling=1, radiosity_mode=1):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        o = Opts(device=0, numerics=numerics, sw_threads=sw_threads, undersampling=undersampling,
                 radiosity_mode=radiosity_mode, rcp_table=None, rcp_bits=0, rsqrt_table=None, rsqrt_bits=0)
        devs = (ctypes.c_int * len(devices))(*devices)
        rc = self._lib.osb200_group_create(ctypes.byref(o), devs, len(devices), ctypes.byref(self._h))
        if rc:
            raise OsbError(f"osb200_group_create failed ({rc}): {self._lib.osb200_last_error().decode()}")
        self.devices = list(devices)

    def _check(self, rc):
        if rc < 0:
            raise OsbError(f"osb200 error {rc}: {self._lib.osb200_last_error().decode()}")
        return rc

    def close(self):
        if getattr(self, "_h", None):
            self._lib.osb200_group_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def size(self):
        return self._lib.osb200_group_size(self._h)

    def map_upload(self, solid, color):
        solid = np.ascontiguousarray(solid, np.uint64)
        color = np.ascontiguousarray(color, np.uint32)
        self._check(self._lib.osb200_group_map_upload(self._h, solid.ctypes.data, color.ctypes.data))

    def map_mark_dirty(self, x, y, z):
        self._check(self._lib.osb200_group_map_mark_dirty(self._h, int(x), int(y), int(z)))

    def map_flush(self, solid, color):
        return self._check(self._lib.osb200_group_map_flush(self._h, solid.ctypes.data, color.ctypes.data))

    def render(self, scenes, w, h, fog_color, fog_dist, want_depth=True, out_bgrx=None, out_depth=None):
        scenes = np.ascontiguousarray(np.atleast_1d(scenes))
        assert scenes.dtype == SCENE_DTYPE
        n = len(scenes)
        fb = np.empty((n, h, w), np.uint32) if out_bgrx is None else out_bgrx
        dp = (np.empty((n, h, w), np.float32) if out_depth is None else out_depth) if want_depth else None
        fc = np.asarray(fog_color, np.float32)
        self._check(self._lib.osb200_group_render(self._h, scenes.ctypes.data, n, w, h, fc.ctypes.data, float(fog_dist),
                                                  _ptr(fb), _ptr(dp)))
        return fb, dp

    def render_gather(self, scenes, w, h, fog_color, fog_dist, root, d_bgrx, d_depth=None):
        """frames of all shards end on device `root` (torch CUDA tensors or raw device pointers)"""
        scenes = np.ascontiguousarray(np.atleast_1d(scenes))
        fc = np.asarray(fog_color, np.float32)
        self._check(self._lib.osb200_group_render_gather(self._h, scenes.ctypes.data, len(scenes), w, h, fc.ctypes.data,
                                                         float(fog_dist), int(root), _ptr(d_bgrx), _ptr(d_depth)))

    def castray2(self, v0, dirs, max_steps, out=None):
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = v0.shape[0]
        if out is None:
            out = np.zeros(n, RAYHIT_DTYPE)
        self._check(self._lib.osb200_group_castray2(self._h, v0.ctypes.data, dirs.ctypes.data, int(max_steps), _ptr(out), n))
        return out
