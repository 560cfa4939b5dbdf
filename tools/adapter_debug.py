"""debug aid: per-frame comparison of the adapter, the plain Context path and SWRenderer on the edit script of
tests/test_gpu_adapter.py::test_edits_change_the_frame"""
import sys, ctypes, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import oracles
import openspades_b200 as osb
from openspades_b200 import workloads
L = ctypes.CDLL(os.path.join(ROOT, 'oracle/_ref/libadapter_sse2.so'))
vp, ci = ctypes.c_void_p, ctypes.c_int
L.adapter_run.argtypes = [vp, ctypes.c_size_t, ci, ci, ci, ci, vp, ci, ci, ci, ctypes.c_uint32, vp, vp, vp, vp]
L.adapter_last_error.restype = ctypes.c_char_p
data = open(oracles.TITLE_VXL, 'rb').read()
vxl = np.frombuffer(data, np.uint8)
for (w, h, edits) in ((256, 256, 300), (256, 256, 40), (320, 200, 300)):
    sc = workloads.title_scene(w, h, 68.0)
    scenes = np.ascontiguousarray(np.stack([sc, sc, sc]))
    n = 3
    a = np.zeros((n, h, w), np.uint32); b = np.zeros((n, h, w), np.uint32); info = np.zeros(4, np.int32)
    ed = np.zeros((2000, 6), np.int32)
    rc = L.adapter_run(vxl.ctypes.data, len(vxl), w, h, 2, 4, scenes.ctypes.data, n, edits, 0, 5, a.ctypes.data, b.ctypes.data,
                       info.ctypes.data, ed.ctypes.data)
    print('case', w, h, edits, 'rc', rc, L.adapter_last_error(), info)
    port = oracles.Port()
    solid, color = port.vxl_load(data)
    with osb.Context(0, osb.NUM_SSE2) as ctx:
        ctx.map_upload(solid, color)
        k = 0
        for f in range(n):
            while k < info[2] and ed[k, 0] == f:
                _, x, y, z, s_, c_ = [int(v) for v in ed[k]]
                if z >= 0:
                    bit = np.uint64(1) << np.uint64(z)
                    if s_:
                        solid[x, y] |= bit; color[x, y, z] = np.uint32(c_ & 0xffffffff)
                    else:
                        solid[x, y] &= ~bit
                    ctx.map_mark_dirty(x, y, z)
                k += 1
            nf = ctx.map_flush(solid, color)
            fb, dp = ctx.render_scene(scenes[f], w, h, (0.8, 1.0, 1.0), 128.0, None, None)
            fb = np.asarray(fb).reshape(h, w)
            rgba = (0xff000000 | (fb & 0xff00) | ((fb & 0xff) << 16) | ((fb & 0xff0000) >> 16))[::-1]
            d_ab = (a[f] != b[f]); d_cb = (rgba != b[f])
            print(' frame', f, 'flushed', nf, 'adapter==SW', 1 - d_ab.mean(), 'context==SW', 1 - d_cb.mean(),
                  'adapter==context', (a[f] == rgba).mean())
            if d_ab.any():
                ys, xs = np.nonzero(d_ab)
                print('   adapter diff rows', ys.min(), ys.max(), 'cols', xs.min(), xs.max(), 'count', d_ab.sum(),
                      'sample', hex(int(a[f][ys[0], xs[0]])), hex(int(b[f][ys[0], xs[0]])))
