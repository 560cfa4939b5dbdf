"""Loaders for the checkers: oracle/liboracle.so (C restatement) and oracle/_ref/liboref_*.so
(the unmodified reference compiled in place).  TEST-SIDE ONLY -- the product never imports this."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
TITLE_VXL = os.path.join(ROOT, "tests", "golden", "maps", "Title.vxl")

SCENE_DTYPE = np.dtype([("origin", "<f4", 3), ("axis", "<f4", (3, 3)), ("fovX", "<f4"), ("fovY", "<f4"),
                        ("zNear", "<f4"), ("zFar", "<f4")])
MODEL_PARAM_DTYPE = np.dtype([("matrix", "<f4", 16), ("customColor", "<f4", 3)])
LIGHT_DTYPE = np.dtype([("origin", "<f4", 3), ("radius", "<f4"), ("color", "<f4", 3)])
ORC_INSTANCE_DTYPE = np.dtype({"names": ["model", "matrix", "customColor"], "formats": ["<u8", ("<f4", 16), ("<f4", 3)],
                               "offsets": [0, 8, 72], "itemsize": 88})
RAYHIT_DTYPE = np.dtype([("hit", "u1"), ("startSolid", "u1"), ("pad", "u1", 2), ("hitPos", "<f4", 3),
                         ("hitBlock", "<i4", 3), ("normal", "<i4", 3)])

vp = ctypes.c_void_p
ci = ctypes.c_int


def build_port():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "port"])
    return os.path.join(ORACLE_DIR, "liboracle.so")


def build_ref():
    """Only possible where /root/reference exists (this container); the GPU box uses the prebuilt files."""
    if os.path.isdir("/root/reference/Sources"):
        subprocess.check_call(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref"])


def ref_available(variant="sse2"):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", f"liboref_{variant}.so")) or os.path.isdir(
        "/root/reference/Sources")


class Port:
    """oracle/liboracle.so"""

    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_port()
        L = self.L = ctypes.CDLL(path)
        L.orc_vxl_load.argtypes = [vp, ctypes.c_size_t, vp, vp]
        L.orc_vxl_save.argtypes = [vp, vp, vp, ctypes.c_size_t]
        L.orc_vxl_save.restype = ctypes.c_longlong
        L.orc_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
        L.orc_castray2.restype = None
        L.orc_build_rle.argtypes = [vp, ci, ci, vp]
        L.orc_render.argtypes = [vp, vp, vp, ci, ci, vp, ctypes.c_float, ci, ci, ci, vp, vp]
        L.orc_last_lines.argtypes = [ctypes.POINTER(ci), ctypes.POINTER(vp), ctypes.POINTER(vp)]
        L.orc_shadow_bitmap.argtypes = [vp, vp, vp]
        L.orc_shadow_bitmap.restype = None
        L.orc_shadow_mark.argtypes = [vp, ci, ci, ci]
        L.orc_shadow_mark.restype = None
        L.orc_shadow_update.argtypes = [vp, vp, vp, vp, vp]
        L.orc_radiosity_invalidate.argtypes = [vp, ci, ci, ci]
        L.orc_radiosity_invalidate.restype = None
        L.orc_radiosity_evaluate.argtypes = [vp, ci, ci, ci, vp]
        L.orc_radiosity_evaluate.restype = None
        L.orc_radiosity_encode.argtypes = [vp, ci]
        L.orc_radiosity_encode.restype = ctypes.c_uint32
        L.orc_radiosity_box.argtypes = [vp] + [ci] * 7 + [vp, vp, ci]
        L.orc_radiosity_box.restype = None
        L.orc_castray.argtypes = [vp, vp, vp, ctypes.c_float, vp, ctypes.c_size_t, ci]
        L.orc_castray.restype = None
        L.orc_ao_evaluate.argtypes = [vp, vp, ci, ci, ci]
        L.orc_ao_evaluate.restype = ctypes.c_float
        L.orc_ao_evaluate_box.argtypes = [vp, vp] + [ci] * 6 + [vp, ci]
        L.orc_ao_evaluate_box.restype = None
        L.orc_ao_state_init.argtypes = [vp]
        L.orc_ao_state_init.restype = None
        L.orc_ao_invalidate.argtypes = [vp, ci, ci, ci]
        L.orc_ao_invalidate.restype = None
        L.orc_ao_update_chunk.argtypes = [vp, vp, vp, ci, ci, ci, vp]
        L.orc_ao_update_chunk.restype = None
        L.orc_links_create.restype = vp
        L.orc_links_free.argtypes = [vp]
        L.orc_links_free.restype = None
        L.orc_links_rebuild.argtypes = [vp, vp]
        L.orc_links_rebuild.restype = None
        L.orc_links_add_block.argtypes = [vp, vp, vp, ci, ci, ci, ctypes.c_uint32]
        L.orc_links_add_block.restype = None
        L.orc_links_remove_blocks.argtypes = [vp, vp, vp, ci, vp, ci]
        L.orc_links_linked.argtypes = [vp, vp]
        L.orc_links_linked.restype = None
        L.orc_minimap.argtypes = [vp, vp, vp]
        L.orc_minimap.restype = None
        L.orc_model_create.argtypes = [ci, ci, ci, vp, vp, vp]
        L.orc_model_create.restype = vp
        L.orc_model_free.argtypes = [vp]
        L.orc_model_free.restype = None
        L.orc_model_render_data.argtypes = [vp, vp, ci, vp, ci]
        L.orc_render_scene.argtypes = [vp, vp, vp, ci, ci, vp, ctypes.c_float, ci, ci, ci, vp, ci, vp, ci, vp, vp]

    # ---- floating blocks + minimap (orc_links.c) ----
    def clusterize_blocks(self, cells):
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros(len(cells), np.int32)
        self.L.orc_clusterize_blocks.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        return self.L.orc_clusterize_blocks(cells.ctypes.data, len(cells), out.ctypes.data), out

    def links_create(self, solid):
        w = self.L.orc_links_create()
        self.L.orc_links_rebuild(w, solid.ctypes.data)
        return w

    def links_free(self, w):
        self.L.orc_links_free(w)

    def links_add_block(self, w, solid, color, x, y, z, col):
        self.L.orc_links_add_block(w, solid.ctypes.data, color.ctypes.data, x, y, z, col)

    def links_remove_blocks(self, w, solid, cells):
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros((1 << 20, 3), np.int16)
        n = self.L.orc_links_remove_blocks(w, solid.ctypes.data, cells.ctypes.data, len(cells), out.ctypes.data, len(out))
        assert 0 <= n <= len(out)
        return out[:n].copy()

    def links_linked(self, w):
        out = np.zeros((512, 512), np.uint64)
        self.L.orc_links_linked(w, out.ctypes.data)
        return out

    def minimap(self, solid, color):
        out = np.zeros((512, 512), np.uint32)
        self.L.orc_minimap(solid.ctypes.data, color.ctypes.data, out.ctypes.data)
        return out

    # ---- voxel models + dynamic lights (orc_models.c) ----
    def model_create(self, w, h, d, origin, solid, colors):
        o = np.asarray(origin, np.float32)
        solid = np.ascontiguousarray(solid, np.uint64)
        colors = np.ascontiguousarray(colors, np.uint32)
        assert solid.size == w * h and colors.size == w * h * d
        m = self.L.orc_model_create(w, h, d, o.ctypes.data, solid.ctypes.data, colors.ctypes.data)
        if not m:
            raise ValueError("orc_model_create failed")
        return m

    def model_free(self, m):
        self.L.orc_model_free(m)

    def model_render_data(self, m, ncols):
        n = self.L.orc_model_render_data(m, None, 0, None, 0)
        data = np.zeros(n, np.uint32)
        addr = np.zeros(ncols, np.uint32)
        self.L.orc_model_render_data(m, data.ctypes.data, n, addr.ctypes.data, ncols)
        return data, addr

    def render_scene(self, solid, color, scene, w, h, fog_color, fog_dist, numerics, models=(), params=None, lights=None,
                     threads=4, under=1, depth=None):
        """EndScene with models (handles + MODEL_PARAM_DTYPE[n]) and point lights (LIGHT_DTYPE[m])."""
        fb = np.zeros((h, w), np.uint32)
        dp = np.zeros((h, w), np.float32) if depth is None else depth
        sc = np.ascontiguousarray(scene)
        fc = np.asarray(fog_color, np.float32)
        inst = np.zeros(len(models), ORC_INSTANCE_DTYPE)
        for i, m in enumerate(models):
            inst[i]["model"] = m
            inst[i]["matrix"] = params[i]["matrix"]
            inst[i]["customColor"] = params[i]["customColor"]
        lights = np.zeros(0, LIGHT_DTYPE) if lights is None else np.ascontiguousarray(lights)
        rc = self.L.orc_render_scene(None if solid is None else solid.ctypes.data,
                                     None if color is None else color.ctypes.data, sc.ctypes.data, w, h, fc.ctypes.data,
                                     fog_dist, numerics, threads, under, inst.ctypes.data, len(inst), lights.ctypes.data,
                                     len(lights), fb.ctypes.data, dp.ctypes.data)
        if rc:
            raise ValueError("orc_render_scene failed")
        return fb, dp

    def vxl_load(self, data: bytes):
        solid = np.zeros((512, 512), np.uint64)
        color = np.zeros((512, 512, 64), np.uint32)
        buf = np.frombuffer(data, np.uint8)
        rc = self.L.orc_vxl_load(buf.ctypes.data, len(data), solid.ctypes.data, color.ctypes.data)
        if rc:
            raise ValueError("truncated VXL")
        return solid, color

    def vxl_save(self, solid, color) -> bytes:
        n = self.L.orc_vxl_save(solid.ctypes.data, color.ctypes.data, None, 0)
        buf = np.zeros(n, np.uint8)
        self.L.orc_vxl_save(solid.ctypes.data, color.ctypes.data, buf.ctypes.data, n)
        return buf.tobytes()

    def castray2(self, solid, v0, dirs, max_steps, threads=1):
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros(len(v0), RAYHIT_DTYPE)
        self.L.orc_castray2(solid.ctypes.data, v0.ctypes.data, dirs.ctypes.data, max_steps, out.ctypes.data,
                            len(v0), threads)
        return out

    def render(self, solid, color, scene, w, h, fog_color, fog_dist, numerics, threads=4, under=1, depth=None):
        fb = np.zeros((h, w), np.uint32)
        dp = np.zeros((h, w), np.float32) if depth is None else depth
        sc = np.ascontiguousarray(scene)
        fc = np.asarray(fog_color, np.float32)
        rc = self.L.orc_render(solid.ctypes.data, color.ctypes.data, sc.ctypes.data, w, h, fc.ctypes.data,
                               fog_dist, numerics, threads, under, fb.ctypes.data, dp.ctypes.data)
        if rc:
            raise ValueError("orc_render failed")
        return fb, dp

    def last_lines(self):
        lr = ci()
        pc, pd = vp(), vp()
        n = self.L.orc_last_lines(ctypes.byref(lr), ctypes.byref(pc), ctypes.byref(pd))
        comb = np.ctypeslib.as_array(ctypes.cast(pc, ctypes.POINTER(ctypes.c_uint32)), shape=(n, lr.value)).copy()
        dep = np.ctypeslib.as_array(ctypes.cast(pd, ctypes.POINTER(ctypes.c_float)), shape=(n, lr.value)).copy()
        return comb, dep

    def build_rle(self, solid, x, y):
        out = np.zeros(400, np.int8)
        n = self.L.orc_build_rle(solid.ctypes.data, x, y, out.ctypes.data)
        return out[:n].copy()

    # ---- shadow bitmap + radiosity ----
    def shadow_bitmap(self, solid, color):
        out = np.zeros((512, 512), np.uint32)  # [y][x]
        self.L.orc_shadow_bitmap(solid.ctypes.data, color.ctypes.data, out.ctypes.data)
        return out

    def shadow_mark(self, words, x, y, z):
        self.L.orc_shadow_mark(words.ctypes.data, x, y, z)

    def shadow_update(self, solid, color, bitmap, words, rad_state=None):
        return self.L.orc_shadow_update(solid.ctypes.data, color.ctypes.data, bitmap.ctypes.data, words.ctypes.data,
                                        None if rad_state is None else rad_state.ctypes.data)

    def radiosity_evaluate(self, bitmap, x, y, z):
        out = np.zeros(12, np.float32)
        self.L.orc_radiosity_evaluate(bitmap.ctypes.data, x, y, z, out.ctypes.data)
        return out

    def radiosity_encode(self, v, mode):
        v = np.ascontiguousarray(v, np.float32)
        return int(self.L.orc_radiosity_encode(v.ctypes.data, mode))

    def radiosity_box(self, bitmap, lo, hi, mode, threads=8, want_f32=True, want_enc=True):
        """Box [lo, hi] inclusive; returns (f32[nz,ny,nx,12], enc[nz,ny,nx,4])."""
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        f32 = np.zeros((nz, ny, nx, 12), np.float32) if want_f32 else None
        enc = np.zeros((nz, ny, nx, 4), np.uint32) if want_enc else None
        self.L.orc_radiosity_box(bitmap.ctypes.data, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], mode,
                                 None if f32 is None else f32.ctypes.data, None if enc is None else enc.ctypes.data,
                                 threads)
        return f32, enc


    # ---- GameMap::CastRay + ambient occlusion ----
    def castray(self, solid, v0, dirs, length, threads=1):
        """(n,4) int32: hit, hitBlock.xyz"""
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros((len(v0), 4), np.int32)
        self.L.orc_castray(solid.ctypes.data, v0.ctypes.data, dirs.ctypes.data, float(length), out.ctypes.data,
                           len(v0), threads)
        return out

    def ao_evaluate(self, solid, rays, x, y, z):
        rays = np.ascontiguousarray(rays, np.float32)
        return float(self.L.orc_ao_evaluate(solid.ctypes.data, rays.ctypes.data, x, y, z))

    def ao_evaluate_box(self, solid, rays, lo, hi, threads=8):
        rays = np.ascontiguousarray(rays, np.float32)
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        out = np.zeros((nz, ny, nx), np.float32)
        self.L.orc_ao_evaluate_box(solid.ctypes.data, rays.ctypes.data, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2],
                                   out.ctypes.data, threads)
        return out

    def ao_state_init(self):
        st = np.zeros((4096, 7), np.int32)
        self.L.orc_ao_state_init(st.ctypes.data)
        return st

    def ao_invalidate(self, state, x, y, z):
        self.L.orc_ao_invalidate(state.ctypes.data, x, y, z)

    def ao_update_chunk(self, solid, rays, state, cx, cy, cz, data=None):
        """data: (16,16,16,2) float32 [z][y][x][ao, air]; a fresh chunk is all ones"""
        rays = np.ascontiguousarray(rays, np.float32)
        if data is None:
            data = np.ones((16, 16, 16, 2), np.float32)
        self.L.orc_ao_update_chunk(solid.ctypes.data, rays.ctypes.data, state.ctypes.data, cx, cy, cz, data.ctypes.data)
        return data


class Ref:
    """oracle/_ref/liboref_{variant}.so -- the reference's own code."""

    def __init__(self, variant="sse2"):
        path = os.path.join(ORACLE_DIR, "_ref", f"liboref_{variant}.so")
        if not os.path.exists(path):
            build_ref()
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.L = ctypes.CDLL(path)
        self.variant = variant
        L.oref_variant.restype = ctypes.c_char_p
        assert L.oref_variant().decode() == variant
        L.oref_last_error.restype = ctypes.c_char_p
        for f in ("oref_map_load_vxl", "oref_map_load_vxl_mem", "oref_map_from_arrays", "oref_renderer_create"):
            getattr(L, f).restype = vp
        L.oref_map_load_vxl_mem.argtypes = [vp, ctypes.c_size_t]
        L.oref_map_from_arrays.argtypes = [vp, vp]
        L.oref_map_save_vxl.argtypes = [vp, vp, ctypes.c_size_t]
        L.oref_map_save_vxl.restype = ctypes.c_longlong
        L.oref_map_free.argtypes = [vp]
        L.oref_map_get_arrays.argtypes = [vp, vp, vp]
        L.oref_map_set.argtypes = [vp, ci, ci, ci, ci, ctypes.c_uint32]
        L.oref_wrapper_create.argtypes = [vp]
        L.oref_wrapper_create.restype = vp
        L.oref_wrapper_free.argtypes = [vp]
        L.oref_wrapper_free.restype = None
        L.oref_wrapper_add_block.argtypes = [vp, ci, ci, ci, ctypes.c_uint32]
        L.oref_wrapper_add_block.restype = None
        L.oref_wrapper_remove_blocks.argtypes = [vp, vp, ci, vp, ci]
        L.oref_wrapper_linked.argtypes = [vp, vp]
        L.oref_wrapper_linked.restype = None
        L.oref_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
        L.oref_castray.argtypes = [vp, vp, vp, ctypes.c_float, vp, ctypes.c_size_t, ci]
        L.oref_castray.restype = None
        L.oref_renderer_create.argtypes = [ci] * 4
        L.oref_renderer_free.argtypes = [vp]
        L.oref_renderer_level.argtypes = [vp]
        L.oref_renderer_set_threads.argtypes = [ci]
        L.oref_renderer_set_map.argtypes = [vp, vp]
        L.oref_renderer_set_fog.argtypes = [vp] + [ctypes.c_float] * 4
        L.oref_renderer_minimap.argtypes = [vp, vp]
        L.oref_renderer_render.argtypes = [vp, vp, vp, vp]
        L.oref_renderer_rle.argtypes = [vp, ci, ci, vp, ci]
        L.oref_renderer_lines.argtypes = [vp, ctypes.POINTER(ci)]
        L.oref_renderer_line.argtypes = [vp, ci, vp, vp, vp, vp]

    def map_load_vxl(self, data: bytes):
        m = self.L.oref_map_load_vxl_mem(data, len(data))
        if not m:
            raise ValueError(self.L.oref_last_error().decode())
        return m

    def map_from_arrays(self, solid, color):
        return self.L.oref_map_from_arrays(solid.ctypes.data, color.ctypes.data)

    def map_save_vxl(self, m) -> bytes:
        """GameMap::Save"""
        n = self.L.oref_map_save_vxl(m, None, 0)
        if n < 0:
            raise ValueError(self.L.oref_last_error().decode())
        buf = np.zeros(n, np.uint8)
        self.L.oref_map_save_vxl(m, buf.ctypes.data, n)
        return buf.tobytes()

    def map_arrays(self, m):
        solid = np.zeros((512, 512), np.uint64)
        color = np.zeros((512, 512, 64), np.uint32)
        self.L.oref_map_get_arrays(m, solid.ctypes.data, color.ctypes.data)
        return solid, color

    # GameMapWrapper (floating-block detection); create = ctor + Rebuild
    def wrapper_create(self, m):
        w = self.L.oref_wrapper_create(m)
        if not w:
            raise RuntimeError(self.L.oref_last_error().decode())
        return w

    def wrapper_free(self, w):
        self.L.oref_wrapper_free(w)

    # World::ClusterizeBlocks, the unmodified static function of Client/World.cpp (oref_world.cpp)
    def clusterize_blocks(self, cells):
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.full(len(cells), -1, np.int32)
        self.L.oref_clusterize_blocks.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        return self.L.oref_clusterize_blocks(cells.ctypes.data, len(cells), out.ctypes.data), out

    def wrapper_add_block(self, w, x, y, z, col):
        self.L.oref_wrapper_add_block(w, x, y, z, int(col))

    def wrapper_remove_blocks(self, w, cells):
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros((1 << 20, 3), np.int16)
        n = self.L.oref_wrapper_remove_blocks(w, cells.ctypes.data, len(cells), out.ctypes.data, len(out))
        if n < 0:
            raise RuntimeError(self.L.oref_last_error().decode())
        return out[:n].copy()

    def wrapper_linked(self, w):
        out = np.zeros((512, 512), np.uint64)
        self.L.oref_wrapper_linked(w, out.ctypes.data)
        return out

    def map_set(self, m, x, y, z, solid, color):
        self.L.oref_map_set(m, x, y, z, int(solid), int(color))

    def map_free(self, m):
        self.L.oref_map_free(m)

    def castray2(self, m, v0, dirs, max_steps, threads=1):
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros(len(v0), RAYHIT_DTYPE)
        self.L.oref_castray2(m, v0.ctypes.data, dirs.ctypes.data, max_steps, out.ctypes.data, len(v0), threads)
        return out

    def castray(self, m, v0, dirs, length, threads=1):
        """GameMap::CastRay: (n,4) int32 hit, hitBlock.xyz"""
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros((len(v0), 4), np.int32)
        self.L.oref_castray(m, v0.ctypes.data, dirs.ctypes.data, float(length), out.ctypes.data, len(v0), threads)
        return out


class RefAmbient:
    """GLAmbientShadowRenderer CPU half of the reference (oracle/ref_harness/oref_gl.cpp)."""

    def __init__(self, ref: Ref, m):
        L = self.L = ref.L
        L.oref_ao_create.restype = vp
        L.oref_ao_create.argtypes = [vp]
        L.oref_ao_free.argtypes = [vp]
        L.oref_ao_get_rays.argtypes = [vp, vp]
        L.oref_ao_set_rays.argtypes = [vp, vp]
        L.oref_ao_evaluate.argtypes = [vp, ci, ci, ci]
        L.oref_ao_evaluate.restype = ctypes.c_float
        L.oref_ao_evaluate_box.argtypes = [vp] + [ci] * 6 + [vp, ci]
        L.oref_ao_map_changed.argtypes = [vp, ci, ci, ci]
        L.oref_ao_chunk_info.argtypes = [vp, ci, ci, ci, vp]
        L.oref_ao_num_dirty.argtypes = [vp]
        L.oref_ao_update_chunk.argtypes = [vp, ci, ci, ci]
        L.oref_ao_update_all.argtypes = [vp, ci]
        L.oref_ao_read_chunk.argtypes = [vp, ci, ci, ci, vp]
        self.h = L.oref_ao_create(m)

    def close(self):
        if self.h:
            self.L.oref_ao_free(self.h)
            self.h = None

    def rays(self):
        out = np.zeros((16, 3), np.float32)
        self.L.oref_ao_get_rays(self.h, out.ctypes.data)
        return out

    def set_rays(self, rays):
        rays = np.ascontiguousarray(rays, np.float32)
        self.L.oref_ao_set_rays(self.h, rays.ctypes.data)

    def evaluate(self, x, y, z):
        return float(self.L.oref_ao_evaluate(self.h, x, y, z))

    def evaluate_box(self, lo, hi, threads=8):
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        out = np.zeros((nz, ny, nx), np.float32)
        self.L.oref_ao_evaluate_box(self.h, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], out.ctypes.data, threads)
        return out

    def map_changed(self, x, y, z):
        self.L.oref_ao_map_changed(self.h, x, y, z)

    def chunk_info(self, cx, cy, cz):
        out = np.zeros(7, np.int32)
        self.L.oref_ao_chunk_info(self.h, cx, cy, cz, out.ctypes.data)
        return out

    def chunk_state(self):
        out = np.zeros((4096, 7), np.int32)
        for cx in range(32):
            for cy in range(32):
                for cz in range(4):
                    out[(cx + cy * 32) * 4 + cz] = self.chunk_info(cx, cy, cz)
        return out

    def num_dirty(self):
        return self.L.oref_ao_num_dirty(self.h)

    def update_chunk(self, cx, cy, cz):
        self.L.oref_ao_update_chunk(self.h, cx, cy, cz)

    def update_all(self, threads=8):
        return self.L.oref_ao_update_all(self.h, threads)

    def read_chunk(self, cx, cy, cz):
        """(16,16,16,2) float32 [z][y][x][ao, air]"""
        out = np.zeros((16, 16, 16, 2), np.float32)
        self.L.oref_ao_read_chunk(self.h, cx, cy, cz, out.ctypes.data)
        return out


class RefRadiosity:
    """GLMapShadowRenderer + GLRadiosityRenderer CPU halves of the reference (oracle/ref_harness/oref_gl.cpp)."""

    def __init__(self, ref: Ref, m, mode=1):
        L = self.L = ref.L
        L.oref_rad_create.restype = vp
        L.oref_rad_create.argtypes = [vp, ci]
        L.oref_rad_free.argtypes = [vp]
        L.oref_rad_map_changed.argtypes = [vp, ci, ci, ci]
        L.oref_rad_shadow_update.argtypes = [vp]
        L.oref_rad_shadow_bitmap.argtypes = [vp, vp]
        L.oref_rad_generate_pixel.argtypes = [vp, ci, ci]
        L.oref_rad_generate_pixel.restype = ctypes.c_uint32
        L.oref_rad_evaluate.argtypes = [vp, ci, ci, ci, vp]
        L.oref_rad_evaluate_box.argtypes = [vp] + [ci] * 6 + [vp, ci]
        L.oref_rad_encode.argtypes = [vp, vp]
        L.oref_rad_encode.restype = ctypes.c_uint32
        L.oref_rad_chunk_info.argtypes = [vp, ci, ci, ci, vp]
        L.oref_rad_num_dirty.argtypes = [vp]
        L.oref_rad_update_chunk.argtypes = [vp, ci, ci, ci]
        L.oref_rad_update_all.argtypes = [vp, ci]
        L.oref_rad_read_chunk.argtypes = [vp, ci, ci, ci, vp, vp, vp, vp]
        self.h = L.oref_rad_create(m, mode)

    def close(self):
        if self.h:
            self.L.oref_rad_free(self.h)
            self.h = None

    def map_changed(self, x, y, z):
        self.L.oref_rad_map_changed(self.h, x, y, z)

    def shadow_update(self):
        self.L.oref_rad_shadow_update(self.h)

    def shadow_bitmap(self):
        out = np.zeros((512, 512), np.uint32)
        self.L.oref_rad_shadow_bitmap(self.h, out.ctypes.data)
        return out

    def evaluate(self, x, y, z):
        out = np.zeros(12, np.float32)
        self.L.oref_rad_evaluate(self.h, x, y, z, out.ctypes.data)
        return out

    def evaluate_box(self, lo, hi, threads=8):
        nx, ny, nz = hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1
        out = np.zeros((nz, ny, nx, 12), np.float32)
        self.L.oref_rad_evaluate_box(self.h, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], out.ctypes.data, threads)
        return out

    def encode(self, v):
        v = np.ascontiguousarray(v, np.float32)
        return int(self.L.oref_rad_encode(self.h, v.ctypes.data))

    def chunk_info(self, cx, cy, cz):
        out = np.zeros(7, np.int32)
        self.L.oref_rad_chunk_info(self.h, cx, cy, cz, out.ctypes.data)
        return out

    def chunk_state(self):
        """All 4096 chunks x 7 ints in the oracle port's order (cx + cy*32)*4 + cz."""
        out = np.zeros((4096, 7), np.int32)
        for cx in range(32):
            for cy in range(32):
                for cz in range(4):
                    out[(cx + cy * 32) * 4 + cz] = self.chunk_info(cx, cy, cz)
        return out

    def num_dirty(self):
        return self.L.oref_rad_num_dirty(self.h)

    def update_chunk(self, cx, cy, cz):
        self.L.oref_rad_update_chunk(self.h, cx, cy, cz)

    def update_all(self, threads=8):
        return self.L.oref_rad_update_all(self.h, threads)

    def read_chunk(self, cx, cy, cz):
        """(16,16,16,4) uint32 [z][y][x][flat,x,y,z]"""
        a = [np.zeros((16, 16, 16), np.uint32) for _ in range(4)]
        self.L.oref_rad_read_chunk(self.h, cx, cy, cz, *[v.ctypes.data for v in a])
        return np.stack(a, axis=-1)


class RefVoxelModel:
    """spades::VoxelModel of the reference (Core/VoxelModel.cpp) -- KV6 loader or raw arrays."""

    def __init__(self, ref: Ref, kv6: bytes = None, arrays=None):
        L = self.L = ref.L
        L.oref_vmodel_load_kv6.restype = vp
        L.oref_vmodel_load_kv6.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.oref_vmodel_from_arrays.restype = vp
        L.oref_vmodel_from_arrays.argtypes = [ci, ci, ci, vp, vp, vp]
        L.oref_vmodel_free.argtypes = [vp]
        L.oref_vmodel_info.argtypes = [vp, vp, vp]
        L.oref_vmodel_arrays.argtypes = [vp, vp, vp]
        if kv6 is not None:
            self.h = L.oref_vmodel_load_kv6(kv6, len(kv6))
        else:
            w, h, d, origin, solid, colors = arrays
            o = np.asarray(origin, np.float32)
            solid = np.ascontiguousarray(solid, np.uint64)
            colors = np.ascontiguousarray(colors, np.uint32)
            self.h = L.oref_vmodel_from_arrays(w, h, d, o.ctypes.data, solid.ctypes.data, colors.ctypes.data)
        if not self.h:
            raise ValueError(L.oref_last_error().decode())
        dims = np.zeros(3, np.int32)
        org = np.zeros(3, np.float32)
        L.oref_vmodel_info(self.h, dims.ctypes.data, org.ctypes.data)
        self.w, self.hgt, self.d = (int(v) for v in dims)
        self.origin = org

    def arrays(self):
        """solid uint64[h][w] (index x + y*w), colors uint32[h][w][d]"""
        solid = np.zeros((self.hgt, self.w), np.uint64)
        colors = np.zeros((self.hgt, self.w, self.d), np.uint32)
        self.L.oref_vmodel_arrays(self.h, solid.ctypes.data, colors.ctypes.data)
        return solid, colors

    def close(self):
        if self.h:
            self.L.oref_vmodel_free(self.h)
            self.h = None


class RefRenderer:
    def __init__(self, ref: Ref, w, h, level=-1, threads=4, under=0):
        """threads / under set the process-wide cvars r_swNumThreads / r_swUndersampling (read on every frame)"""
        self.ref, self.w, self.h = ref, w, h
        L = ref.L
        L.oref_renderer_set_undersampling(under)
        L.oref_model_create.restype = vp
        L.oref_model_create.argtypes = [vp, vp]
        L.oref_model_free.argtypes = [vp]
        L.oref_model_render_data.argtypes = [vp, vp, ci, vp, ci]
        L.oref_renderer_render_scene.argtypes = [vp, vp, vp, vp, ci, vp, ci, vp, vp]
        self.r = ref.L.oref_renderer_create(w, h, level, threads)
        if not self.r:
            raise RuntimeError(ref.L.oref_last_error().decode())

    def set_map(self, m):
        self.ref.L.oref_renderer_set_map(self.r, m)

    def set_fog(self, color, dist):
        self.ref.L.oref_renderer_set_fog(self.r, color[0], color[1], color[2], dist)

    def render(self, scene):
        fb = np.zeros((self.h, self.w), np.uint32)
        dp = np.zeros((self.h, self.w), np.float32)
        sc = np.ascontiguousarray(scene)
        rc = self.ref.L.oref_renderer_render(self.r, sc.ctypes.data, fb.ctypes.data, dp.ctypes.data)
        if rc:
            raise RuntimeError(self.ref.L.oref_last_error().decode())
        return fb, dp

    def minimap(self):
        out = np.zeros((512, 512), np.uint32)
        if self.ref.L.oref_renderer_minimap(self.r, out.ctypes.data):
            raise RuntimeError(self.ref.L.oref_last_error().decode())
        return out

    def create_model(self, vmodel: "RefVoxelModel"):
        m = self.ref.L.oref_model_create(self.r, vmodel.h)
        if not m:
            raise RuntimeError(self.ref.L.oref_last_error().decode())
        return m

    def free_model(self, m):
        self.ref.L.oref_model_free(m)

    def model_render_data(self, m, ncols):
        n = self.ref.L.oref_model_render_data(m, None, 0, None, 0)
        data = np.zeros(n, np.uint32)
        addr = np.zeros(ncols, np.uint32)
        self.ref.L.oref_model_render_data(m, data.ctypes.data, n, addr.ctypes.data, ncols)
        return data, addr

    def render_scene(self, scene, models=(), params=None, lights=None):
        """StartScene + RenderModel* + AddLight* + EndScene; params: MODEL_PARAM_DTYPE[n], lights: LIGHT_DTYPE[m]"""
        fb = np.zeros((self.h, self.w), np.uint32)
        dp = np.zeros((self.h, self.w), np.float32)
        sc = np.ascontiguousarray(scene)
        arr = (vp * max(1, len(models)))(*models)
        params = np.zeros(0, MODEL_PARAM_DTYPE) if params is None else np.ascontiguousarray(params)
        lights = np.zeros(0, LIGHT_DTYPE) if lights is None else np.ascontiguousarray(lights)
        rc = self.ref.L.oref_renderer_render_scene(self.r, sc.ctypes.data, arr, params.ctypes.data, len(models),
                                                   lights.ctypes.data, len(lights), fb.ctypes.data, dp.ctypes.data)
        if rc:
            raise RuntimeError(self.ref.L.oref_last_error().decode())
        return fb, dp

    def rle(self, x, y):
        out = np.zeros(400, np.int8)
        n = self.ref.L.oref_renderer_rle(self.r, x, y, out.ctypes.data, 400)
        return out[:n].copy()

    def lines(self):
        lr = ci()
        n = self.ref.L.oref_renderer_lines(self.r, ctypes.byref(lr))
        return n, lr.value

    def line(self, i, line_res):
        comb = np.zeros(line_res, np.uint32)
        dep = np.zeros(line_res, np.float32)
        misc = np.zeros(4, np.float32)
        misci = np.zeros(2, np.int32)
        n = self.ref.L.oref_renderer_line(self.r, i, comb.ctypes.data, dep.ctypes.data, misc.ctypes.data,
                                          misci.ctypes.data)
        return n, comb, dep, misc, misci

    def close(self):
        if self.r:
            self.ref.L.oref_renderer_free(self.r)
            self.r = None
