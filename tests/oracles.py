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
        L.orc_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
        L.orc_castray2.restype = None
        L.orc_build_rle.argtypes = [vp, ci, ci, vp]
        L.orc_render.argtypes = [vp, vp, vp, ci, ci, vp, ctypes.c_float, ci, ci, ci, vp, vp]
        L.orc_last_lines.argtypes = [ctypes.POINTER(ci), ctypes.POINTER(vp), ctypes.POINTER(vp)]

    def vxl_load(self, data: bytes):
        solid = np.zeros((512, 512), np.uint64)
        color = np.zeros((512, 512, 64), np.uint32)
        buf = np.frombuffer(data, np.uint8)
        rc = self.L.orc_vxl_load(buf.ctypes.data, len(data), solid.ctypes.data, color.ctypes.data)
        if rc:
            raise ValueError("truncated VXL")
        return solid, color

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
        L.oref_map_free.argtypes = [vp]
        L.oref_map_get_arrays.argtypes = [vp, vp, vp]
        L.oref_map_set.argtypes = [vp, ci, ci, ci, ci, ctypes.c_uint32]
        L.oref_castray2.argtypes = [vp, vp, vp, ci, vp, ctypes.c_size_t, ci]
        L.oref_renderer_create.argtypes = [ci] * 4
        L.oref_renderer_free.argtypes = [vp]
        L.oref_renderer_level.argtypes = [vp]
        L.oref_renderer_set_threads.argtypes = [ci]
        L.oref_renderer_set_map.argtypes = [vp, vp]
        L.oref_renderer_set_fog.argtypes = [vp] + [ctypes.c_float] * 4
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

    def map_arrays(self, m):
        solid = np.zeros((512, 512), np.uint64)
        color = np.zeros((512, 512, 64), np.uint32)
        self.L.oref_map_get_arrays(m, solid.ctypes.data, color.ctypes.data)
        return solid, color

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


class RefRenderer:
    def __init__(self, ref: Ref, w, h, level=-1, threads=4):
        self.ref, self.w, self.h = ref, w, h
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
