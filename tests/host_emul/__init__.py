"""TEST INFRASTRUCTURE: the product's per-ray CUDA source (csrc/rt_device.cuh + csrc/rt_shade.cuh) compiled as plain
C++ (emul.cpp) and driven from Python, so the parity tests can run the same functions on the CPU against the oracle.
Nothing in the package imports this; it is not a CPU fallback."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "triangles_sdf_cpu_raytracing_b200", "csrc")
OUT = os.path.join(HERE, "_build", "libemul.so")
SOURCES = [os.path.join(HERE, "emul.cpp"), os.path.join(CSRC, "bvh_builder.cpp")]
DEPS = SOURCES + [os.path.join(CSRC, n) for n in ("rt_device.cuh", "rt_shade.cuh", "rt_scene.h", "bvh_builder.h")]


def cuda_include() -> str | None:
    for d in (os.environ.get("CUDA_HOME"), "/usr/local/cuda"):
        if d and os.path.exists(os.path.join(d, "include", "cuda_runtime.h")):
            return os.path.join(d, "include")
    return None


def build(force: bool = False) -> str:
    """g++ -ffp-contract=off: one IEEE operation per operator, as nvcc -fmad=false gives the exact flavour."""
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in DEPS):
        return OUT
    inc = cuda_include()
    if inc is None:
        raise FileNotFoundError("cuda_runtime.h not found (the device headers use float4 / make_float4 from it)")
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-Wno-unknown-pragmas",
           "-I", inc, "-I", CSRC, "-o", OUT] + SOURCES
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("host emulation build failed:\n" + r.stderr[-4000:])
    return OUT


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        vp = C.c_void_p
        L.emul_grid.restype = vp
        L.emul_grid.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32]
        L.emul_grid_bricks.restype = vp
        L.emul_grid_bricks.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32]
        L.emul_intersect.restype = C.c_int
        L.emul_intersect.argtypes = [vp, vp, vp, C.c_uint32, C.c_float, C.c_float, C.c_int, C.c_float, vp, vp, vp, vp]
        for fn in (L.emul_intersect_walk, L.emul_intersect_octree_counted):
            fn.restype = C.c_int
            fn.argtypes = [vp, vp, vp, C.c_uint32, C.c_float, C.c_float, vp, vp, vp, vp]
        L.emul_work_map.restype = C.c_int
        L.emul_work_map.argtypes = [vp, vp, vp]
        L.emul_octree.restype = vp
        L.emul_octree.argtypes = [vp, C.c_uint32]
        L.emul_mesh.restype = vp
        L.emul_mesh.argtypes = [vp, C.c_uint32, vp, C.c_uint32, C.c_int]
        L.emul_mesh_refbvh.restype = vp
        L.emul_mesh_refbvh.argtypes = [vp, C.c_uint32, vp, C.c_uint32, vp, C.c_uint32, C.c_uint32, vp]
        L.emul_free.argtypes = [vp]
        L.emul_render.restype = C.c_int
        L.emul_render.argtypes = [vp, vp, vp, vp, vp, vp]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


COUNTER_NAMES = ("rays", "hits", "grid_samples", "oct_inner", "oct_leaf", "oct_leaf_samples", "bvh_inner", "bvh_leaf",
                 "bvh_tris")


class EmulScene:
    def __init__(self, handle, keep=()):
        if not handle:
            raise ValueError("host emulation: scene construction failed")
        self._h = C.c_void_p(handle)
        self._keep = keep

    def __del__(self):
        if getattr(self, "_h", None):
            lib().emul_free(self._h)
            self._h = None

    @staticmethod
    def grid(values, bricks=False) -> "EmulScene":
        v = np.ascontiguousarray(values, np.float32)
        return EmulScene((lib().emul_grid_bricks if bricks else lib().emul_grid)(_p(v), *v.shape))

    @staticmethod
    def octree(nodes36) -> "EmulScene":
        r = np.ascontiguousarray(nodes36, np.uint8).reshape(-1, 36)
        return EmulScene(lib().emul_octree(_p(r), r.shape[0]))

    @staticmethod
    def mesh(vpos4, indices, brute=None) -> "EmulScene":
        v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
        i = np.ascontiguousarray(indices, np.uint32).reshape(-1)
        # rt_scene_create_mesh's rules: <= 8 triangles are the reference's one-leaf tree (BRUTE_ROOT_LEAF = 2) unless a
        # tree is asked for; RT_MESH_AUTO (brute=None) builds a tree beyond that; RT_MESH_BRUTE (brute=True) is
        # BRUTE_BOXED = 1 there
        n_tri = i.size // 3
        mode = 0 if brute is False else 2 if n_tri <= 8 else 1 if brute else 0
        return EmulScene(lib().emul_mesh(_p(v), v.shape[0], _p(i), i.size, mode))

    @staticmethod
    def mesh_refbvh(vpos4, indices_bvh, nodes, prim_of_permuted=None) -> "EmulScene":
        v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
        i = np.ascontiguousarray(indices_bvh, np.uint32).reshape(-1)
        n = np.ascontiguousarray(nodes)
        nb = n.view(np.uint8).reshape(n.shape[0], -1)
        p = None if prim_of_permuted is None else np.ascontiguousarray(prim_of_permuted, np.uint32)
        return EmulScene(lib().emul_mesh_refbvh(_p(v), v.shape[0], _p(i), i.size, _p(nb), nb.shape[0], nb.shape[1], _p(p)))

    def render(self, frame, color=None, t=None, want_prim=True, count=True):
        """frame: the package's RtFrame (rt.make_frame).  -> (color [H,W] u32, t [H,W] f32, prim, counters dict).
        count=False runs the kernels' plain flavour (no workload counters), the one the timed launches use."""
        H, W = frame.height, frame.width
        color = np.zeros((H, W), np.uint32) if color is None else np.ascontiguousarray(color, np.uint32).copy()
        t = np.full((H, W), np.inf, np.float32) if t is None else np.ascontiguousarray(t, np.float32).copy()
        prim = np.full((H, W), 0xFFFFFFFF, np.uint32) if want_prim else None
        cnt = np.zeros(9, np.uint64)
        rc = lib().emul_render(self._h, C.byref(frame), _p(color), _p(t), _p(prim), _p(cnt) if count else None)
        assert rc == 0
        return color, t, prim, dict(zip(COUNTER_NAMES, (int(x) for x in cnt)))

    def intersect(self, origins, dirs, t_near=0.01, t_far=100.0, plane_y=None):
        """IScene::intersect per ray -> (hit bool[n], t f32[n], normal f32[n,3], prim u32[n])."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        hit = np.zeros(n, np.uint8)
        t = np.zeros(n, np.float32)
        nrm = np.zeros((n, 3), np.float32)
        prim = np.zeros(n, np.uint32)
        rc = lib().emul_intersect(self._h, _p(o), _p(d), n, t_near, t_far, int(plane_y is not None),
                                  float(plane_y or 0.0), _p(hit), _p(t), _p(nrm), _p(prim))
        assert rc == 0
        return hit.astype(bool), t, nrm, prim

    def work_map(self, frame):
        """Per-pixel work of the primary rays -> uint32 [H, W, 4]: inner visits, leaf visits, SDF samples, triangle tests."""
        w = np.zeros((frame.height, frame.width, 4), np.uint32)
        assert lib().emul_work_map(self._h, C.byref(frame), _p(w)) == 0
        return w

    def octree_rays(self, origins, dirs, t_near=0.01, t_far=100.0, walk=False):
        """Octree scenes: oct_intersect (walk=False) or the resumable OctWalk form driven to completion (walk=True)
        -> (hit, t, normal, (inner visits, leaf visits, leaf samples))."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        hit, t, nrm, cnt = np.zeros(n, np.uint8), np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(3, np.uint64)
        fn = lib().emul_intersect_walk if walk else lib().emul_intersect_octree_counted
        assert fn(self._h, _p(o), _p(d), n, t_near, t_far, _p(hit), _p(t), _p(nrm), _p(cnt)) == 0
        return hit.astype(bool), t, nrm, tuple(int(x) for x in cnt)
