"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libref_oracle*.so (oracle/ref_driver.cpp).

Only tests/, bench.py's cpu_baseline / --impl reference legs and __graft_entry__.smoke() may import
this module.  It runs the reference's own C++ (Renderer::draw, IScene::intersect, BVHBuilder,
loadSDFGrid/loadSDFOctree) compiled by oracle/Makefile; the product never routes through it.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KIND_GRID, KIND_OCTREE, KIND_MESH, KIND_UNION = 0, 1, 2, 3
MODE_NORMAL, MODE_LAMBERT, MODE_COLOR = 0, 1, 2  # raytracing.hpp:99
MISS_PRIM = 0xFFFFFFFF


class RefCam(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("target", C.c_float * 3), ("up", C.c_float * 3),
                ("lockUp", C.c_int32), ("nRotate", C.c_int32), ("dx", C.c_float), ("dy", C.c_float)]


@dataclass
class CamSpec:
    """Camera description; the default is the viewer's start camera (main.cpp:89-91)."""
    eye: tuple = (0.0, 0.0, 2.5)
    target: tuple = (0.0, 0.0, 0.0)
    up: tuple = (0.0, 1.0, 0.0)
    lock_up: bool = True
    n_rotate: int = 0
    dx: float = 0.0
    dy: float = 0.0

    def c(self) -> RefCam:
        r = RefCam()
        r.eye[:] = self.eye
        r.target[:] = self.target
        r.up[:] = self.up
        r.lockUp = int(self.lock_up)
        r.nRotate = int(self.n_rotate)
        r.dx = self.dx
        r.dy = self.dy
        return r


import contextlib
import sys


@contextlib.contextmanager
def quiet_stdout():
    """BVHBuilder::perform prints its build time to std::cout (triangles_raytracing.cpp:255-257); keep it out of
    the caller's stdout (bench.py prints exactly one JSON line there)."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    try:
        os.dup2(devnull, 1)
        yield
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def lib_path(contract: bool = False) -> str:
    return os.path.join(_HERE, "_ref", "libref_oracle_contract.so" if contract else "libref_oracle.so")


def available(contract: bool = False) -> bool:
    return os.path.exists(lib_path(contract))


class RefOracle:
    """One loaded flavour of the reference library (canonical -ffp-contract=off, or 'contract')."""

    def __init__(self, contract: bool = False):
        path = lib_path(contract)
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
        self.contract = contract
        L = self.L = C.CDLL(path)
        vp = C.c_void_p
        f32p, u32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_max_threads.restype = C.c_int
        for name, args in [
            ("ref_scene_grid_file", [C.c_char_p]),
            ("ref_scene_grid", [f32p, C.c_uint32, C.c_uint32, C.c_uint32]),
            ("ref_scene_octree_file", [C.c_char_p]),
            ("ref_scene_octree", [vp, C.c_uint32]),
            ("ref_scene_mesh", [f32p, C.c_uint32, u32p, C.c_uint32, C.c_int]),
            ("ref_scene_mesh_obj", [C.c_char_p, C.c_int]),
            ("ref_scene_with_plane", [vp, C.c_float]),
        ]:
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = vp
        L.ref_scene_model_min_y.argtypes = [vp]
        L.ref_scene_model_min_y.restype = C.c_float
        L.ref_scene_kind.argtypes = [vp]
        L.ref_scene_free.argtypes = [vp]
        L.ref_mesh_counts.argtypes = [vp, u32p, u32p, u32p, u32p]
        L.ref_mesh_export.argtypes = [vp, f32p, u32p, vp, u32p]
        L.ref_camera.argtypes = [C.POINTER(RefCam), f32p, f32p, f32p]
        L.ref_projection.argtypes = [C.c_float] * 4 + [f32p, f32p]
        L.ref_render.argtypes = [vp, C.POINTER(RefCam), C.c_uint32, C.c_uint32, f32p, C.c_int, f32p,
                                 C.c_int, C.c_int, u32p, f32p]
        L.ref_render.restype = C.c_float
        L.ref_render_timed.argtypes = [vp, C.POINTER(RefCam), C.c_uint32, C.c_uint32, f32p, C.c_int, f32p,
                                       C.c_int, C.c_int, C.c_int, f32p]
        L.ref_render_timed.restype = C.c_float
        L.ref_render_rows.argtypes = [vp, C.POINTER(RefCam), C.c_uint32, C.c_uint32, f32p, C.c_uint32,
                                      C.c_uint32, u32p, f32p]
        L.ref_render_rows.restype = C.c_float
        L.ref_render_prim.argtypes = [vp, C.POINTER(RefCam), C.c_uint32, C.c_uint32, f32p, f32p, f32p, u32p]
        L.ref_intersect.argtypes = [vp, f32p, f32p, C.c_float, C.c_float, f32p, f32p]
        L.ref_intersect.restype = C.c_int
        L.ref_grid_sdf.argtypes = [vp, f32p]
        L.ref_grid_sdf.restype = C.c_float
        L.ref_grid_normal.argtypes = [vp, f32p, f32p]
        L.ref_bbox_intersection.argtypes = [f32p, f32p, f32p, f32p, C.c_float, C.c_float, f32p, f32p,
                                            C.POINTER(C.c_int)]
        L.ref_eye_ray.argtypes = [C.c_uint32] * 4 + [f32p, f32p, f32p]
        L.ref_color_pack.argtypes = [f32p]
        L.ref_color_pack.restype = C.c_uint32
        L.ref_sort8.argtypes = [f32p, C.POINTER(C.c_uint64)]
        L.ref_divide_box_8.argtypes = [f32p, f32p]
        L.ref_intersect_box_8.argtypes = [f32p, f32p, f32p, C.c_float, C.c_float, f32p]
        L.ref_intersect_1_ray_8_triangles.argtypes = [f32p, f32p, f32p, u32p, f32p, f32p]

    # ---- scenes
    def set_threads(self, n: int):
        self.L.ref_set_threads(int(n))

    def max_threads(self) -> int:
        return self.L.ref_max_threads()

    def grid(self, values: np.ndarray) -> "RefSceneHandle":
        v = np.ascontiguousarray(values, dtype=np.float32)
        sx, sy, sz = v.shape
        return RefSceneHandle(self, self.L.ref_scene_grid(_fp(v), sx, sy, sz))

    def grid_file(self, path: str) -> "RefSceneHandle":
        return RefSceneHandle(self, self.L.ref_scene_grid_file(path.encode()))

    def octree(self, nodes36: np.ndarray) -> "RefSceneHandle":
        raw = np.ascontiguousarray(nodes36).view(np.uint8).reshape(-1)
        assert raw.size % 36 == 0
        return RefSceneHandle(self, self.L.ref_scene_octree(raw.ctypes.data_as(C.c_void_p), raw.size // 36))

    def octree_file(self, path: str) -> "RefSceneHandle":
        return RefSceneHandle(self, self.L.ref_scene_octree_file(path.encode()))

    def mesh(self, vpos4: np.ndarray, indices: np.ndarray, scale: bool = False) -> "RefSceneHandle":
        v = np.ascontiguousarray(vpos4, dtype=np.float32).reshape(-1, 4)
        i = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        with quiet_stdout():
            h = self.L.ref_scene_mesh(_fp(v), v.shape[0], _up(i), i.size, int(scale))
        return RefSceneHandle(self, h)

    def mesh_obj(self, path: str, scale: bool = True) -> "RefSceneHandle":
        with quiet_stdout():
            h = self.L.ref_scene_mesh_obj(path.encode(), int(scale))
        if not h:
            raise RuntimeError(f"reference could not load {path}")
        return RefSceneHandle(self, h)

    # ---- camera / projection
    def camera(self, cam: CamSpec):
        pos = np.zeros(3, np.float32)
        view = np.zeros(16, np.float32)
        view_inv = np.zeros(16, np.float32)
        c = cam.c()
        self.L.ref_camera(C.byref(c), _fp(pos), _fp(view), _fp(view_inv))
        return pos, view, view_inv

    def projection(self, W: int, H: int, fovy=45.0, z_near=0.01, z_far=100.0):
        """main.cpp:198-201. Returns (proj, projInv) as 16 floats, column-major."""
        proj = np.zeros(16, np.float32)
        inv = np.zeros(16, np.float32)
        aspect = np.float32(W) / np.float32(H)
        self.L.ref_projection(fovy, float(aspect), z_near, z_far, _fp(proj), _fp(inv))
        return proj, inv

    # ---- primitive KATs
    def bbox_intersection(self, bmin, bmax, o, inv_dir, t_min, t_max):
        a = [np.ascontiguousarray(x, np.float32) for x in (bmin, bmax, o, inv_dir)]
        t1, t2, face = C.c_float(), C.c_float(), C.c_int()
        self.L.ref_bbox_intersection(_fp(a[0]), _fp(a[1]), _fp(a[2]), _fp(a[3]), t_min, t_max,
                                     C.byref(t1), C.byref(t2), C.byref(face))
        return np.float32(t1.value), np.float32(t2.value), face.value

    def eye_ray(self, x, y, W, H, proj_inv, view_inv):
        out = np.zeros(3, np.float32)
        self.L.ref_eye_ray(x, y, W, H, _fp(np.ascontiguousarray(proj_inv, np.float32)),
                           _fp(np.ascontiguousarray(view_inv, np.float32)), _fp(out))
        return out

    def color_pack(self, c4) -> int:
        return int(self.L.ref_color_pack(_fp(np.ascontiguousarray(c4, np.float32))))

    def sort8(self, t):
        t = np.array(t, np.float32)
        ch = np.arange(8, dtype=np.uint64)
        self.L.ref_sort8(_fp(t), ch.ctypes.data_as(C.POINTER(C.c_uint64)))
        return t, ch.astype(np.int64)

    def divide_box_8(self, box6):
        out = np.zeros(48, np.float32)
        self.L.ref_divide_box_8(_fp(np.ascontiguousarray(box6, np.float32)), _fp(out))
        return out.reshape(6, 8)

    def intersect_box_8(self, boxes6x8, o, inv_dir, t_near, t_far):
        out = np.zeros(8, np.float32)
        b = np.ascontiguousarray(boxes6x8, np.float32).reshape(48)
        self.L.ref_intersect_box_8(_fp(b), _fp(np.ascontiguousarray(o, np.float32)),
                                   _fp(np.ascontiguousarray(inv_dir, np.float32)), t_near, t_far, _fp(out))
        return out

    def intersect_1_ray_8_triangles(self, tris, o, d):
        """tris: [8 triangles, 3 vertices, 3 coords] -> (hitten[8], t[8], normal[8,3])."""
        tr = np.ascontiguousarray(tris, np.float32).reshape(8, 3, 3)
        soa = np.ascontiguousarray(tr.transpose(1, 2, 0)).reshape(72)  # v?.{x,y,z}[8]
        hit = np.zeros(8, np.uint32)
        t = np.zeros(8, np.float32)
        n = np.zeros(24, np.float32)
        self.L.ref_intersect_1_ray_8_triangles(_fp(soa), _fp(np.ascontiguousarray(o, np.float32)),
                                               _fp(np.ascontiguousarray(d, np.float32)), _up(hit), _fp(t), _fp(n))
        return hit, t, n.reshape(3, 8).T.copy()


@dataclass
class RefSceneHandle:
    lib: RefOracle
    h: int
    _children: list = field(default_factory=list)

    def __post_init__(self):
        if not self.h:
            raise RuntimeError("reference scene construction failed")

    def __del__(self):
        try:
            if self.h:
                self.lib.L.ref_scene_free(self.h)
                self.h = None
        except Exception:
            pass

    @property
    def kind(self) -> int:
        return self.lib.L.ref_scene_kind(self.h)

    @property
    def model_min_y(self) -> float:
        return float(self.lib.L.ref_scene_model_min_y(self.h))

    def with_plane(self, plane_y: float) -> "RefSceneHandle":
        s = RefSceneHandle(self.lib, self.lib.L.ref_scene_with_plane(self.h, plane_y))
        s._children.append(self)  # keep the base alive
        return s

    def export_mesh(self):
        """-> dict(vpos4 [nV,4] f32, indices [nIdx] u32 in BVH order, nodes [nNodes, nodeBytes] u8,
        prim_of_permuted [nTri] u32)."""
        nv, ni, nn, nb = (C.c_uint32() for _ in range(4))
        self.lib.L.ref_mesh_counts(self.h, C.byref(nv), C.byref(ni), C.byref(nn), C.byref(nb))
        v = np.zeros((nv.value, 4), np.float32)
        i = np.zeros(ni.value, np.uint32)
        n = np.zeros((nn.value, nb.value), np.uint8)
        p = np.zeros(ni.value // 3, np.uint32)
        self.lib.L.ref_mesh_export(self.h, _fp(v), _up(i), n.ctypes.data_as(C.c_void_p), _up(p))
        return dict(vpos4=v, indices=i, nodes=n, prim_of_permuted=p)

    def render(self, cam: CamSpec, W: int, H: int, proj_inv=None, mode=MODE_NORMAL, light=(2.0, 2.0, 2.0),
               shadows=False, reflections=False, color=None, t=None):
        """Renderer::draw into (color, t). Fresh cleared buffers unless given. Returns (color, t, ms)."""
        if proj_inv is None:
            proj_inv = self.lib.projection(W, H)[1]
        pi = np.ascontiguousarray(proj_inv, np.float32)
        color = np.zeros((H, W), np.uint32) if color is None else np.ascontiguousarray(color, np.uint32).copy()
        t = np.full((H, W), np.inf, np.float32) if t is None else np.ascontiguousarray(t, np.float32).copy()
        li = np.ascontiguousarray(light, np.float32)
        c = cam.c()
        ms = self.lib.L.ref_render(self.h, C.byref(c), W, H, _fp(pi), mode, _fp(li), int(shadows),
                                   int(reflections), _up(color), _fp(t))
        return color, t, float(ms)

    def render_timed(self, cam: CamSpec, W: int, H: int, frames=5, proj_inv=None, mode=MODE_NORMAL,
                     light=(2.0, 2.0, 2.0), shadows=False, reflections=False):
        if proj_inv is None:
            proj_inv = self.lib.projection(W, H)[1]
        pi = np.ascontiguousarray(proj_inv, np.float32)
        li = np.ascontiguousarray(light, np.float32)
        each = np.zeros(frames, np.float32)
        c = cam.c()
        best = self.lib.L.ref_render_timed(self.h, C.byref(c), W, H, _fp(pi), mode, _fp(li), int(shadows),
                                           int(reflections), frames, _fp(each))
        return float(best), each

    def render_rows(self, cam: CamSpec, W: int, H: int, y0: int, y1: int, proj_inv=None):
        if proj_inv is None:
            proj_inv = self.lib.projection(W, H)[1]
        pi = np.ascontiguousarray(proj_inv, np.float32)
        color = np.zeros((H, W), np.uint32)
        t = np.full((H, W), np.inf, np.float32)
        c = cam.c()
        ms = self.lib.L.ref_render_rows(self.h, C.byref(c), W, H, _fp(pi), y0, y1, _up(color), _fp(t))
        return color, t, float(ms)

    def render_prim(self, cam: CamSpec, W: int, H: int, proj_inv=None):
        if proj_inv is None:
            proj_inv = self.lib.projection(W, H)[1]
        pi = np.ascontiguousarray(proj_inv, np.float32)
        t = np.zeros((H, W), np.float32)
        n = np.zeros((H, W, 3), np.float32)
        prim = np.zeros((H, W), np.uint32)
        c = cam.c()
        self.lib.L.ref_render_prim(self.h, C.byref(c), W, H, _fp(pi), _fp(t), _fp(n), _up(prim))
        return t, n, prim

    def intersect(self, o, d, t_near=0.01, t_far=100.0):
        o = np.ascontiguousarray(o, np.float32)
        d = np.ascontiguousarray(d, np.float32)
        t = C.c_float()
        n = np.zeros(3, np.float32)
        hit = self.lib.L.ref_intersect(self.h, _fp(o), _fp(d), t_near, t_far, C.byref(t), _fp(n))
        return bool(hit), np.float32(t.value), n

    def grid_sdf(self, p) -> np.float32:
        return np.float32(self.lib.L.ref_grid_sdf(self.h, _fp(np.ascontiguousarray(p, np.float32))))

    def grid_normal(self, p):
        out = np.zeros(3, np.float32)
        self.lib.L.ref_grid_normal(self.h, _fp(np.ascontiguousarray(p, np.float32)), _fp(out))
        return out
