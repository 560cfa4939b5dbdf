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
