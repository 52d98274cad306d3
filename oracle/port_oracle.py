"""TEST INFRASTRUCTURE — ctypes binding of oracle/libport_oracle.so (oracle/port_oracle.c).

The plain-C restatement of the reference's primary-ray path. It needs nothing but gcc, so it is the
checker that is always available (also on the GPU box, where /root/reference does not exist).
Only tests/, bench.py's cpu_baseline leg and __graft_entry__.smoke() may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libport_oracle.so")

PRIM_MISS = 0xFFFFFFFF
PRIM_PLANE = 0xFFFFFFFE
MODE_NORMAL, MODE_LAMBERT, MODE_COLOR = 0, 1, 2


class PortScene(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("values", C.c_void_p), ("sx", C.c_uint32), ("sy", C.c_uint32), ("sz", C.c_uint32),
        ("nodes36", C.c_void_p), ("nOctNodes", C.c_uint32),
        ("vpos4", C.c_void_p), ("indices", C.c_void_p), ("bvhNodes", C.c_void_p),
        ("nBvhNodes", C.c_uint32), ("bvhNodeBytes", C.c_uint32), ("primOfPermuted", C.c_void_p),
        ("hasPlane", C.c_int32), ("planeY", C.c_float),
        ("bboxMin", C.c_float * 3), ("bboxMax", C.c_float * 3),
    ]


class PortFrame(C.Structure):
    _fields_ = [
        ("rayPos", C.c_float * 3), ("viewInv", C.c_float * 16), ("projInv", C.c_float * 16),
        ("W", C.c_uint32), ("H", C.c_uint32), ("mode", C.c_int32), ("light", C.c_float * 3),
        ("shadows", C.c_int32), ("reflections", C.c_int32),
    ]


class PortCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("rays", "gridSdf", "octInner", "octLeaf", "octLeafSdf", "bvhInner", "bvhLeaf", "bvhTris", "hits")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "port_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        f32p, u32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
        L.port_render.argtypes = [C.POINTER(PortScene), C.POINTER(PortFrame), C.c_uint32, C.c_uint32, u32p, f32p,
                                  u32p, C.POINTER(PortCounters)]
        L.port_intersect.argtypes = [C.POINTER(PortScene), f32p, f32p, C.c_float, C.c_float, f32p, f32p, u32p]
        L.port_intersect.restype = C.c_int
        L.port_eye_ray.argtypes = [C.POINTER(PortFrame), C.c_uint32, C.c_uint32, f32p]
        L.port_grid_sdf.argtypes = [C.POINTER(PortScene), f32p]
        L.port_grid_sdf.restype = C.c_float
        L.port_grid_normal.argtypes = [C.POINTER(PortScene), f32p, f32p]
        L.port_bbox_intersection.argtypes = [f32p, f32p, f32p, f32p, C.c_float, C.c_float, f32p, f32p,
                                             C.POINTER(C.c_int)]
        L.port_intersect_box_8.argtypes = [f32p, f32p, f32p, C.c_float, C.c_float, f32p]
        L.port_sort8.argtypes = [f32p, u32p]
        L.port_triangle.argtypes = [f32p, f32p, f32p, f32p, f32p]
        L.port_triangle.restype = C.c_int
        L.port_color_pack.argtypes = [f32p]
        L.port_color_pack.restype = C.c_uint32
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


class Scene:
    """Host-memory scene in the reference's own formats. Keeps the numpy arrays alive."""

    def __init__(self):
        self.c = PortScene()
        self._keep = []

    @staticmethod
    def grid(values: np.ndarray) -> "Scene":
        s = Scene()
        v = np.ascontiguousarray(values, np.float32)
        assert v.ndim == 3
        s._keep.append(v)
        s.c.kind = 0
        s.c.values = v.ctypes.data
        s.c.sx, s.c.sy, s.c.sz = v.shape
        return s

    @staticmethod
    def octree(nodes36: np.ndarray) -> "Scene":
        s = Scene()
        raw = np.ascontiguousarray(nodes36).view(np.uint8).reshape(-1)
        assert raw.size % 36 == 0
        s._keep.append(raw)
        s.c.kind = 1
        s.c.nodes36 = raw.ctypes.data
        s.c.nOctNodes = raw.size // 36
        return s

    @staticmethod
    def mesh(vpos4, indices, nodes, prim_of_permuted=None) -> "Scene":
        s = Scene()
        v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
        i = np.ascontiguousarray(indices, np.uint32).reshape(-1)
        n = np.ascontiguousarray(nodes, np.uint8)
        assert n.ndim == 2
        s._keep += [v, i, n]
        s.c.kind = 2
        s.c.vpos4 = v.ctypes.data
        s.c.indices = i.ctypes.data
        s.c.bvhNodes = n.ctypes.data
        s.c.nBvhNodes = n.shape[0]
        s.c.bvhNodeBytes = n.shape[1]
        if prim_of_permuted is not None:
            p = np.ascontiguousarray(prim_of_permuted, np.uint32)
            s._keep.append(p)
            s.c.primOfPermuted = p.ctypes.data
        return s

    @staticmethod
    def mesh_brute(vpos4, indices) -> "Scene":
        """Mesh without a BVH: every ray tests every triangle in index order (tree-independent closest hit)."""
        s = Scene()
        v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
        i = np.ascontiguousarray(indices, np.uint32).reshape(-1)
        s._keep += [v, i]
        s.c.kind = 2
        s.c.vpos4 = v.ctypes.data
        s.c.indices = i.ctypes.data
        s.c.bvhNodes = None
        s.c.nBvhNodes = 0
        s.c.bvhNodeBytes = i.size // 3  # triangle count (see port_oracle.c mesh_brute)
        used = v[i]                                      # vertices the triangles reference
        p = (used[:, :3] / used[:, 3:4]).astype(np.float32)  # the division by w of triangles_raytracing.cpp:307-309
        s.c.bboxMin[:] = [float(x) for x in p.min(axis=0)]
        s.c.bboxMax[:] = [float(x) for x in p.max(axis=0)]
        return s

    def with_plane(self, y: float) -> "Scene":
        s = Scene()
        C.memmove(C.byref(s.c), C.byref(self.c), C.sizeof(PortScene))
        s._keep = list(self._keep)
        s.c.hasPlane = 1
        s.c.planeY = y
        return s

    # ---- rendering
    def frame(self, ray_pos, view_inv, proj_inv, W, H, mode=MODE_NORMAL, light=(2.0, 2.0, 2.0), shadows=False,
              reflections=False) -> PortFrame:
        f = PortFrame()
        f.rayPos[:] = [float(x) for x in ray_pos]
        f.viewInv[:] = [float(x) for x in np.asarray(view_inv, np.float32).reshape(16)]
        f.projInv[:] = [float(x) for x in np.asarray(proj_inv, np.float32).reshape(16)]
        f.W, f.H, f.mode = W, H, mode
        f.light[:] = light
        f.shadows, f.reflections = int(shadows), int(reflections)
        return f

    def render(self, ray_pos, view_inv, proj_inv, W, H, mode=MODE_NORMAL, light=(2.0, 2.0, 2.0), shadows=False,
               reflections=False, color=None, t=None, rows=None, want_counters=False):
        """-> (color[H,W] u32, t[H,W] f32, prim[H,W] u32, counters dict or None)."""
        f = self.frame(ray_pos, view_inv, proj_inv, W, H, mode, light, shadows, reflections)
        color = np.zeros((H, W), np.uint32) if color is None else np.ascontiguousarray(color, np.uint32).copy()
        t = np.full((H, W), np.inf, np.float32) if t is None else np.ascontiguousarray(t, np.float32).copy()
        prim = np.full((H, W), PRIM_MISS, np.uint32)
        y0, y1 = rows if rows is not None else (0, H)
        cnt = PortCounters()
        lib().port_render(C.byref(self.c), C.byref(f), y0, y1, _up(color), _fp(t), _up(prim),
                          C.byref(cnt) if want_counters else None)
        return color, t, prim, (cnt.as_dict() if want_counters else None)

    def intersect(self, o, d, t_near=0.01, t_far=100.0):
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        t = C.c_float()
        n = np.zeros(3, np.float32)
        prim = C.c_uint32()
        hit = lib().port_intersect(C.byref(self.c), _fp(o), _fp(d), t_near, t_far, C.byref(t), _fp(n), C.byref(prim))
        return bool(hit), np.float32(t.value), n, int(prim.value)

    def grid_sdf(self, p):
        return np.float32(lib().port_grid_sdf(C.byref(self.c), _fp(np.ascontiguousarray(p, np.float32))))

    def grid_normal(self, p):
        out = np.zeros(3, np.float32)
        lib().port_grid_normal(C.byref(self.c), _fp(np.ascontiguousarray(p, np.float32)), _fp(out))
        return out


# ---- primitive KAT helpers
def bbox_intersection(bmin, bmax, o, inv_dir, t_min, t_max):
    a = [np.ascontiguousarray(x, np.float32) for x in (bmin, bmax, o, inv_dir)]
    t1, t2, face = C.c_float(), C.c_float(), C.c_int()
    lib().port_bbox_intersection(_fp(a[0]), _fp(a[1]), _fp(a[2]), _fp(a[3]), t_min, t_max, C.byref(t1), C.byref(t2),
                                 C.byref(face))
    return np.float32(t1.value), np.float32(t2.value), face.value


def intersect_box_8(boxes6x8, o, inv_dir, t_near, t_far):
    out = np.zeros(8, np.float32)
    b = np.ascontiguousarray(boxes6x8, np.float32).reshape(48)
    lib().port_intersect_box_8(_fp(b), _fp(np.ascontiguousarray(o, np.float32)),
                               _fp(np.ascontiguousarray(inv_dir, np.float32)), t_near, t_far, _fp(out))
    return out


def sort8(t):
    t = np.array(t, np.float32)
    ch = np.arange(8, dtype=np.uint32)
    lib().port_sort8(_fp(t), _up(ch))
    return t, ch.astype(np.int64)


def triangle(o, d, tri9):
    t = C.c_float()
    n = np.zeros(3, np.float32)
    hit = lib().port_triangle(_fp(np.ascontiguousarray(o, np.float32)), _fp(np.ascontiguousarray(d, np.float32)),
                              _fp(np.ascontiguousarray(tri9, np.float32).reshape(9)), C.byref(t), _fp(n))
    return bool(hit), np.float32(t.value), n


def eye_ray(ray_pos, view_inv, proj_inv, W, H, x, y):
    f = Scene().frame(ray_pos, view_inv, proj_inv, W, H)
    out = np.zeros(3, np.float32)
    lib().port_eye_ray(C.byref(f), x, y, _fp(out))
    return out


def color_pack(c4) -> int:
    return int(lib().port_color_pack(_fp(np.ascontiguousarray(c4, np.float32))))
