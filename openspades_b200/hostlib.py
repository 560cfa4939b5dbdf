"""ctypes view of the C++ host layer (openspades_b200/host/: VoxelMap, B200Renderer) via its C wrappers."""
import ctypes
import os

import numpy as np

from .binding import OsbError, load_library

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "lib", "libosb200_host.so")
_hlib = None
vp = ctypes.c_void_p


def load_host_library():
    global _hlib
    if _hlib is not None:
        return _hlib
    load_library()
    if not os.path.exists(HOST_LIB_PATH):
        raise OsbError(f"{HOST_LIB_PATH} not built: run `python -m openspades_b200.build`")
    L = ctypes.CDLL(HOST_LIB_PATH)
    L.osbh_last_error.restype = ctypes.c_char_p
    for f in ("osbh_map_new", "osbh_map_load", "osbh_map_load_file", "osbh_map_solid", "osbh_map_color"):
        getattr(L, f).restype = vp
    L.osbh_map_load.argtypes = [vp, ctypes.c_size_t]
    L.osbh_map_load_file.argtypes = [ctypes.c_char_p]
    L.osbh_map_free.argtypes = [vp]
    L.osbh_map_free.restype = None
    L.osbh_map_solid.argtypes = [vp]
    L.osbh_map_color.argtypes = [vp]
    L.osbh_map_set.argtypes = [vp] + [ctypes.c_int] * 4 + [ctypes.c_uint32]
    L.osbh_map_set_arrays.argtypes = [vp, vp, vp]
    _hlib = L
    return L


class VoxelMap:
    """osb200::VoxelMap (host/VoxelMap.h) -- the GameMap-shaped host truth."""

    def __init__(self, handle=None):
        self._L = load_host_library()
        self._h = handle if handle is not None else self._L.osbh_map_new()

    @classmethod
    def load(cls, data: bytes):
        L = load_host_library()
        buf = np.frombuffer(data, np.uint8)
        h = L.osbh_map_load(buf.ctypes.data, len(data))
        if not h:
            raise OsbError(L.osbh_last_error().decode())
        return cls(h)

    @classmethod
    def load_file(cls, path):
        L = load_host_library()
        h = L.osbh_map_load_file(os.fsencode(path))
        if not h:
            raise OsbError(L.osbh_last_error().decode())
        return cls(h)

    @classmethod
    def from_arrays(cls, solid, color):
        m = cls()
        solid = np.ascontiguousarray(solid, np.uint64)
        color = np.ascontiguousarray(color, np.uint32)
        m._L.osbh_map_set_arrays(m._h, solid.ctypes.data, color.ctypes.data)
        return m

    @property
    def handle(self):
        return self._h

    @property
    def solid(self):
        """Live view of solid[x][y] (no copy)."""
        p = self._L.osbh_map_solid(self._h)
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint64)), shape=(512, 512))

    @property
    def color(self):
        p = self._L.osbh_map_color(self._h)
        return np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint32)), shape=(512, 512, 64))

    def set(self, x, y, z, solid, color):
        if self._L.osbh_map_set(self._h, x, y, z, int(solid), int(color)):
            raise OsbError(self._L.osbh_last_error().decode())

    def close(self):
        if self._h:
            self._L.osbh_map_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class B200Renderer:
    """osb200::B200Renderer (host/B200Renderer.h): SWRenderer's surface over the C ABI, on an in-memory port."""

    def __init__(self, w, h, device=0, numerics=2, sw_threads=4, undersampling=1, radiosity_mode=1):
        L = self._L = load_host_library()
        ci, fp = ctypes.c_int, ctypes.c_float
        L.osbh_renderer_new.restype = vp
        L.osbh_renderer_new.argtypes = [ci] * 7
        L.osbh_renderer_context.restype = vp
        for f in ("free", "init", "shutdown", "end_scene", "frame_done", "flip", "swap_count", "add_sprite",
                  "last_flushed", "context"):
            getattr(L, "osbh_renderer_" + f).argtypes = [vp]
        L.osbh_renderer_free.restype = None
        L.osbh_renderer_set_map.argtypes = [vp, vp]
        L.osbh_renderer_load_map.argtypes = [vp, vp, ctypes.c_size_t]
        L.osbh_renderer_load_map.restype = vp
        L.osbh_renderer_save_map.argtypes = [vp, vp, ctypes.c_size_t]
        L.osbh_renderer_save_map.restype = ctypes.c_longlong
        L.osbh_renderer_set_fog.argtypes = [vp, fp, fp, fp, fp]
        L.osbh_renderer_start_scene.argtypes = [vp, vp]
        L.osbh_renderer_read_bitmap.argtypes = [vp, vp]
        L.osbh_renderer_framebuffer.argtypes = [vp, vp]
        L.osbh_renderer_read_depth.argtypes = [vp, vp]
        L.osbh_renderer_screen.argtypes = [vp, ctypes.POINTER(fp), ctypes.POINTER(fp)]
        L.osbh_renderer_update_radiosity.argtypes = [vp, ci, ctypes.POINTER(ci)]
        L.osbh_renderer_update_ambient.argtypes = [vp, ci, ctypes.POINTER(ci)]
        L.osbh_renderer_create_model.argtypes = [vp, ci, ci, ci, vp, vp, vp, ctypes.POINTER(vp)]
        L.osbh_model_free.argtypes = [vp]
        L.osbh_model_free.restype = None
        L.osbh_renderer_render_model.argtypes = [vp, vp, vp, vp]
        L.osbh_renderer_add_light.argtypes = [vp, ci, vp, ctypes.c_float, vp]
        L.osbh_renderer_set_ambient_rays.argtypes = [vp, vp]
        L.osbh_map_castray.argtypes = [vp, vp, vp, vp, ctypes.c_float, vp, ctypes.c_size_t]
        L.osbh_map_castray2.argtypes = [vp, vp, vp, vp, ci, vp, ctypes.c_size_t]
        self.w, self.h = w, h
        self._h = L.osbh_renderer_new(w, h, device, numerics, sw_threads, undersampling, radiosity_mode)
        if not self._h:
            raise OsbError(L.osbh_last_error().decode())

    def _chk(self, rc):
        if rc:
            raise OsbError(self._L.osbh_last_error().decode())

    def init(self):
        self._chk(self._L.osbh_renderer_init(self._h))

    def shutdown(self):
        self._chk(self._L.osbh_renderer_shutdown(self._h))

    def set_game_map(self, m):
        self._map = m  # keep the host map alive while it is attached
        self._chk(self._L.osbh_renderer_set_map(self._h, None if m is None else m.handle))

    def load_game_map(self, data: bytes):
        """B200Renderer::LoadGameMap: GameMap::Load on the device + SetGameMap; returns the VoxelMap."""
        buf = np.frombuffer(data, np.uint8)
        h = self._L.osbh_renderer_load_map(self._h, buf.ctypes.data, len(buf))
        if not h:
            self._map = None
            raise OsbError(self._L.osbh_last_error().decode())
        self._map = VoxelMap(h)
        return self._map

    def save_game_map(self) -> bytes:
        """B200Renderer::SaveGameMap: GameMap::Save of the current map, encoded on the device."""
        n = self._L.osbh_renderer_save_map(self._h, None, 0)
        if n < 0:
            raise OsbError(self._L.osbh_last_error().decode())
        out = np.empty(n, np.uint8)
        if self._L.osbh_renderer_save_map(self._h, out.ctypes.data, n) != n:
            raise OsbError(self._L.osbh_last_error().decode())
        return out.tobytes()

    def set_fog(self, color, dist):
        self._chk(self._L.osbh_renderer_set_fog(self._h, color[0], color[1], color[2], dist))

    def start_scene(self, scene):
        sc = np.ascontiguousarray(scene)
        self._chk(self._L.osbh_renderer_start_scene(self._h, sc.ctypes.data))

    def add_sprite(self):
        self._chk(self._L.osbh_renderer_add_sprite(self._h))

    def end_scene(self):
        self._chk(self._L.osbh_renderer_end_scene(self._h))

    def frame_done(self):
        self._chk(self._L.osbh_renderer_frame_done(self._h))

    def flip(self):
        self._chk(self._L.osbh_renderer_flip(self._h))

    @property
    def swap_count(self):
        return self._L.osbh_renderer_swap_count(self._h)

    @property
    def last_flushed(self):
        return self._L.osbh_renderer_last_flushed(self._h)

    def framebuffer(self):
        out = np.zeros((self.h, self.w), np.uint32)
        self._chk(self._L.osbh_renderer_framebuffer(self._h, out.ctypes.data))
        return out

    def read_bitmap(self):
        out = np.zeros((self.h, self.w), np.uint32)
        self._chk(self._L.osbh_renderer_read_bitmap(self._h, out.ctypes.data))
        return out

    def read_depth(self):
        out = np.zeros((self.h, self.w), np.float32)
        self._chk(self._L.osbh_renderer_read_depth(self._h, out.ctypes.data))
        return out

    def screen(self):
        w, h = ctypes.c_float(), ctypes.c_float()
        self._chk(self._L.osbh_renderer_screen(self._h, ctypes.byref(w), ctypes.byref(h)))
        return w.value, h.value

    def update_radiosity(self, max_chunks=8):
        n = ctypes.c_int()
        self._chk(self._L.osbh_renderer_update_radiosity(self._h, max_chunks, ctypes.byref(n)))
        return n.value

    def create_model(self, w, h, d, origin, solid, colors):
        """B200Renderer::CreateModel from VoxelModel arrays; returns a Model handle (free with free_model)."""
        o = np.asarray(origin, np.float32)
        solid = np.ascontiguousarray(solid, np.uint64)
        colors = np.ascontiguousarray(colors, np.uint32)
        m = ctypes.c_void_p()
        self._chk(self._L.osbh_renderer_create_model(self._h, w, h, d, o.ctypes.data, solid.ctypes.data,
                                                     colors.ctypes.data, ctypes.byref(m)))
        return m.value

    def free_model(self, m):
        self._L.osbh_model_free(m)

    def render_model(self, model, matrix, custom_color=(0, 0, 0)):
        mt = np.ascontiguousarray(matrix, np.float32)
        cc = np.asarray(custom_color, np.float32)
        self._chk(self._L.osbh_renderer_render_model(self._h, model, mt.ctypes.data, cc.ctypes.data))

    def add_light(self, origin, radius, color, light_type=0):
        o = np.asarray(origin, np.float32)
        c = np.asarray(color, np.float32)
        self._chk(self._L.osbh_renderer_add_light(self._h, int(light_type), o.ctypes.data, float(radius), c.ctypes.data))

    def update_ambient(self, max_chunks=8):
        n = ctypes.c_int()
        self._chk(self._L.osbh_renderer_update_ambient(self._h, max_chunks, ctypes.byref(n)))
        return n.value

    def set_ambient_rays(self, rays):
        rays = np.ascontiguousarray(rays, np.float32)
        assert rays.size == 48
        self._chk(self._L.osbh_renderer_set_ambient_rays(self._h, rays.ctypes.data))

    def castray(self, vmap, v0, dirs, length):
        """VoxelMap::CastRayBatch on this renderer's context: (n,4) int32 hit, hitBlock.xyz."""
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros((len(v0), 4), np.int32)
        self._chk(self._L.osbh_map_castray(vmap.handle, self._L.osbh_renderer_context(self._h), v0.ctypes.data,
                                           dirs.ctypes.data, float(length), out.ctypes.data, len(v0)))
        return out

    def castray2(self, vmap, v0, dirs, max_steps):
        """VoxelMap::CastRay2Batch on this renderer's context."""
        from .binding import RAYHIT_DTYPE
        v0 = np.ascontiguousarray(v0, np.float32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.zeros(len(v0), RAYHIT_DTYPE)
        self._chk(self._L.osbh_map_castray2(vmap.handle, self._L.osbh_renderer_context(self._h), v0.ctypes.data,
                                            dirs.ctypes.data, max_steps, out.ctypes.data, len(v0)))
        return out

    def close(self):
        if getattr(self, "_h", None):
            self._L.osbh_renderer_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class VoxelMapWrapper:
    """osb200::VoxelMapWrapper (host/VoxelMapWrapper.h) -- client::GameMapWrapper's surface: Rebuild / AddBlock / RemoveBlocks
    on a VoxelMap whose device copy lives in `renderer`'s context."""

    def __init__(self, vmap: VoxelMap, renderer: "B200Renderer"):
        L = self._L = load_host_library()
        vp = ctypes.c_void_p
        L.osbh_wrapper_new.argtypes = [vp, vp]
        L.osbh_wrapper_new.restype = vp
        L.osbh_wrapper_free.argtypes = [vp]
        L.osbh_wrapper_free.restype = None
        L.osbh_wrapper_rebuild.argtypes = [vp]
        L.osbh_wrapper_add_block.argtypes = [vp] + [ctypes.c_int] * 3 + [ctypes.c_uint32]
        L.osbh_wrapper_remove_blocks.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int]
        L.osbh_wrapper_clusterize.argtypes = [vp, vp, ctypes.c_int, vp, vp, ctypes.c_int]
        L.osbh_renderer_context.restype = vp
        self._map, self._renderer = vmap, renderer  # keep both alive
        self._h = L.osbh_wrapper_new(vmap.handle, L.osbh_renderer_context(renderer._h))

    def _chk(self, rc):
        if rc < 0:
            raise OsbError(self._L.osbh_last_error().decode())
        return rc

    def rebuild(self):
        self._chk(self._L.osbh_wrapper_rebuild(self._h))

    def add_block(self, x, y, z, color):
        self._chk(self._L.osbh_wrapper_add_block(self._h, x, y, z, int(color)))

    def remove_blocks(self, cells):
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros((1 << 16, 3), np.int16)
        n = self._chk(self._L.osbh_wrapper_remove_blocks(self._h, cells.ctypes.data, len(cells), out.ctypes.data, len(out)))
        return out[:min(n, len(out))].copy()

    def clusterize_blocks(self, cells):
        """World::ClusterizeBlocks (World.cpp:154-226): list of (m, 3) int16 arrays, one per 6-connected piece."""
        cells = np.ascontiguousarray(cells, np.int16).reshape(-1, 3)
        out = np.zeros((max(len(cells), 1), 3), np.int16)
        sizes = np.zeros(max(len(cells), 1), np.int32)
        k = self._chk(self._L.osbh_wrapper_clusterize(self._h, cells.ctypes.data, len(cells), out.ctypes.data,
                                                      sizes.ctypes.data, len(sizes)))
        ends = np.cumsum(sizes[:k])
        return [out[e - s:e].copy() for s, e in zip(sizes[:k], ends)]

    def close(self):
        if self._h:
            self._L.osbh_wrapper_free(self._h)
            self._h = None
