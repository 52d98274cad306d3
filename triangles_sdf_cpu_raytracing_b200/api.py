"""Python face of the C ABI (tests and bench.py drive the CUDA path through this; numpy in, numpy out).

Names follow the reference: a scene is an SDFGrid (grid_raytracing.hpp), an SDFOctree (octree_raytracing.hpp) or
a triangle mesh behind a BVH (triangles_raytracing.hpp); `render` is Renderer::draw (raytracing.cpp:67-102) and
`intersect` is IScene::intersect (raytracing.hpp:75-81).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def make_frame(ray_pos, view_inv, proj_inv, W, H, mode=L.SHADE_NORMAL, light=(2.0, 2.0, 2.0), shadows=False,
               reflections=False, plane_y=None, bands=None, t_near=0.0, t_far=0.0, clear=False) -> L.RtFrame:
    """bands = (band_rows, band_count, band_index) selects an interleaved screen-space partition."""
    f = L.RtFrame()
    f.ray_pos[:] = [float(x) for x in np.asarray(ray_pos, np.float32).reshape(3)]
    f.view_inv[:] = [float(x) for x in np.asarray(view_inv, np.float32).reshape(16)]
    f.proj_inv[:] = [float(x) for x in np.asarray(proj_inv, np.float32).reshape(16)]
    f.width, f.height = int(W), int(H)
    if bands is not None:
        f.band_rows, f.band_count, f.band_index = (int(x) for x in bands)
    f.shading_mode = int(mode)
    f.light_pos[:] = [float(x) for x in light]
    f.enable_shadows = int(bool(shadows))
    f.enable_reflections = int(bool(reflections))
    f.has_plane = int(plane_y is not None)
    f.plane_y = float(plane_y) if plane_y is not None else 0.0
    f.t_near, f.t_far = float(t_near), float(t_far)
    f.clear = int(bool(clear))
    return f


class PinnedFrameBuffer:
    """FrameBuffer (raytracing.hpp:9-20) in page-locked host memory: `color` [H,W] uint32 and `t` [H,W] float32 are
    numpy views of rt_pinned_alloc memory, so Scene.render_into stores its tiles straight into them (rt_b200.h,
    rt_render_host).  Keep the object alive while the views are in use."""

    def __init__(self, W: int, H: int):
        self.W, self.H = int(W), int(H)
        n = self.W * self.H * 4
        self._pc, self._pt = C.c_void_p(), C.c_void_p()
        L.check(L.lib().rt_pinned_alloc(n, C.byref(self._pc)))
        L.check(L.lib().rt_pinned_alloc(n, C.byref(self._pt)))
        self.color = np.ctypeslib.as_array((C.c_uint32 * (self.W * self.H)).from_address(self._pc.value)).reshape(H, W)
        self.t = np.ctypeslib.as_array((C.c_float * (self.W * self.H)).from_address(self._pt.value)).reshape(H, W)
        self.clear()

    def clear(self):
        """FrameBuffer::clear (raytracing.hpp:16-19)."""
        self.color[...] = 0
        self.t[...] = np.inf

    def free(self):
        for name in ("_pc", "_pt"):
            p = getattr(self, name, None)
            if p is not None and p.value:
                L.lib().rt_pinned_free(p)
                setattr(self, name, None)
        self.color = self.t = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Scene:
    """An uploaded, immutable scene (rt_scene*)."""

    def __init__(self, handle):
        self._h = handle

    def __del__(self):
        self.destroy()

    def destroy(self):
        if getattr(self, "_h", None):
            L.lib().rt_scene_destroy(self._h)
            self._h = None

    @property
    def handle(self):
        return self._h

    @property
    def kind(self) -> int:
        return L.lib().rt_scene_kind_of(self._h)

    def info(self):
        b = C.c_uint64()
        lo = (C.c_float * 3)()
        hi = (C.c_float * 3)()
        L.check(L.lib().rt_scene_info(self._h, C.byref(b), lo, hi))
        return dict(device_bytes=int(b.value), bbox_min=np.array(lo[:], np.float32), bbox_max=np.array(hi[:], np.float32))

    def transfer_bytes(self):
        """Running (fetched, stored) framebuffer bytes of the in/out calls on pinned / peer buffers (rt_scene_transfer_bytes)."""
        f, s = C.c_uint64(), C.c_uint64()
        L.check(L.lib().rt_scene_transfer_bytes(self._h, C.byref(f), C.byref(s)))
        return int(f.value), int(s.value)

    def replicas(self):
        """Devices that hold a copy of the scene (rt_init)."""
        d = np.zeros(64, np.int32)
        n = L.lib().rt_scene_replicas(self._h, _ptr(d), 64)
        return [int(x) for x in d[:n]]

    def build_ms(self) -> float:
        """Device time of the on-device part of scene construction (octree ingestion, device BVH build)."""
        ms = C.c_float()
        L.check(L.lib().rt_scene_build_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def mesh_stats(self):
        """Mesh scenes: dict(n_nodes, n_tris, device_build_ms) (device_build_ms is 0 for host-built trees)."""
        n, t, ms = C.c_uint32(), C.c_uint32(), C.c_float()
        L.check(L.lib().rt_scene_mesh_stats(self._h, C.byref(n), C.byref(t), C.byref(ms)))
        return dict(n_nodes=int(n.value), n_tris=int(t.value), device_build_ms=float(ms.value))

    # ---- Renderer::draw with host framebuffers
    def render(self, frame: L.RtFrame, color=None, t=None, want_prim=False, want_counters=False):
        """-> dict(color [H,W] u32, t [H,W] f32, prim or None, device_ms, total_ms, counters or None).
        color / t default to a cleared framebuffer (FrameBuffer::clear, raytracing.hpp:16-19)."""
        H, W = frame.height, frame.width
        color = np.zeros((H, W), np.uint32) if color is None else np.ascontiguousarray(color, np.uint32).copy()
        t = np.full((H, W), np.inf, np.float32) if t is None else np.ascontiguousarray(t, np.float32).copy()
        assert color.shape == (H, W) and t.shape == (H, W)
        prim = np.full((H, W), L.PRIM_MISS, np.uint32) if want_prim else None
        cnt = L.RtCounters() if want_counters else None
        dms, tms = C.c_float(), C.c_float()
        L.check(L.lib().rt_render_host(self._h, C.byref(frame), _ptr(color), _ptr(t), _ptr(prim), C.byref(dms),
                                       C.byref(tms), C.byref(cnt) if cnt is not None else None))
        return dict(color=color, t=t, prim=prim, device_ms=dms.value, total_ms=tms.value,
                    counters=cnt.as_dict() if cnt is not None else None)

    def render_into(self, frame: L.RtFrame, fb: PinnedFrameBuffer):
        """Renderer::draw into a page-locked framebuffer, in place -> (device_ms, total_ms)."""
        assert (fb.H, fb.W) == (frame.height, frame.width)
        dms, tms = C.c_float(), C.c_float()
        L.check(L.lib().rt_render_host(self._h, C.byref(frame), fb._pc, fb._pt, None, C.byref(dms), C.byref(tms), None))
        return float(dms.value), float(tms.value)

    # ---- Renderer::draw on device-resident framebuffers (torch tensors or raw pointers), asynchronous
    def render_device(self, frame: L.RtFrame, color_ptr: int, t_in_ptr, t_out_ptr: int, prim_ptr=None,
                      counters_ptr=None, stream: int = 0):
        L.check(L.lib().rt_render_device(self._h, C.byref(frame), C.c_void_p(color_ptr),
                                         C.c_void_p(t_in_ptr) if t_in_ptr else None, C.c_void_p(t_out_ptr),
                                         C.c_void_p(prim_ptr) if prim_ptr else None,
                                         C.c_void_p(counters_ptr) if counters_ptr else None,
                                         C.c_void_p(stream) if stream else None))

    # ---- IScene::intersect for a batch of rays
    def intersect(self, origins, dirs, t_near=0.01, t_far=100.0, plane_y=None):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        hit = np.zeros(n, np.uint8)
        t = np.zeros(n, np.float32)
        nrm = np.zeros((n, 3), np.float32)
        prim = np.zeros(n, np.uint32)
        L.check(L.lib().rt_intersect_host(self._h, _ptr(o), _ptr(d), n, t_near, t_far, int(plane_y is not None),
                                          float(plane_y or 0.0), _ptr(hit), _ptr(t), _ptr(nrm), _ptr(prim)))
        return hit.astype(bool), t, nrm, prim


def _new_scene(fn, *args) -> Scene:
    h = C.c_void_p()
    L.check(fn(*args, C.byref(h)))
    return Scene(h)


def SDFGrid(values: np.ndarray, layout: int = L.GRID_AUTO) -> Scene:
    """values[x, y, z] float32 (grid_raytracing.hpp:13-15)."""
    v = np.ascontiguousarray(values, np.float32)
    if v.ndim != 3:
        raise ValueError("grid values must be a 3-d array [sx, sy, sz]")
    return _new_scene(L.lib().rt_scene_create_grid, _ptr(v), v.shape[0], v.shape[1], v.shape[2], layout)


def SDFOctree(nodes36: np.ndarray) -> Scene:
    """nodes36: the raw 36-byte node records (octree_raytracing.hpp:8-11) as uint8 [n, 36] (or any n*36 bytes)."""
    raw = np.ascontiguousarray(nodes36).view(np.uint8).reshape(-1)
    if raw.size % 36:
        raise ValueError("octree data must be a multiple of 36 bytes")
    return _new_scene(L.lib().rt_scene_create_octree, _ptr(raw), raw.size // 36)


def TriangleMesh(vpos4: np.ndarray, indices: np.ndarray, accel: int = L.MESH_AUTO) -> Scene:
    """BVHBuilder::perform on SimpleMesh::vPos4f / indices (triangles_raytracing.hpp:40)."""
    v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
    i = np.ascontiguousarray(indices, np.uint32).reshape(-1)
    return _new_scene(L.lib().rt_scene_create_mesh, _ptr(v), v.shape[0], _ptr(i), i.size, accel)


def TriangleMeshFromReferenceBVH(vpos4, indices_bvh, nodes, prim_of_permuted=None) -> Scene:
    """Traverse a BVH8 built by the reference's own BVHBuilder (204-byte BVH8Node images)."""
    v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
    i = np.ascontiguousarray(indices_bvh, np.uint32).reshape(-1)
    n = np.ascontiguousarray(nodes, np.uint8)
    p = None if prim_of_permuted is None else np.ascontiguousarray(prim_of_permuted, np.uint32)
    return _new_scene(L.lib().rt_scene_create_mesh_bvh8, _ptr(v), v.shape[0], _ptr(i), i.size, _ptr(n), n.shape[0],
                      n.shape[1], _ptr(p))


def load_grid_file(path: str) -> np.ndarray:
    """.grid: u32 sx, sy, sz then sx*sy*sz f32 (grid_raytracing.cpp:127-134)."""
    raw = np.fromfile(path, dtype=np.uint8)
    if raw.size < 12:
        raise ValueError(f"{path}: truncated header")
    sx, sy, sz = (int(x) for x in raw[:12].view(np.uint32))
    if raw.size != 12 + 4 * sx * sy * sz:
        raise ValueError(f"{path}: size does not match its {sx}x{sy}x{sz} header")
    return raw[12:].view(np.float32).reshape(sx, sy, sz)


def load_octree_file(path: str) -> np.ndarray:
    """.octree: u32 n then n x 36-byte nodes (octree_raytracing.cpp:8-16)."""
    raw = np.fromfile(path, dtype=np.uint8)
    if raw.size < 4:
        raise ValueError(f"{path}: truncated header")
    n = int(raw[:4].view(np.uint32)[0])
    if raw.size != 4 + 36 * n:
        raise ValueError(f"{path}: size does not match its node count {n}")
    return raw[4:].reshape(n, 36)


def init(devices=None):
    """rt_init: scenes created afterwards are replicated on every listed device and ONE render call fans the frame out
    over them (include/rt_b200.h).  None / [] restores single-device scenes on the current device."""
    d = np.ascontiguousarray(list(devices or []), np.int32)
    L.check(L.lib().rt_init(_ptr(d) if d.size else None, int(d.size)))


def set_math_mode(mode: int):
    L.check(L.lib().rt_set_math_mode(mode))


def set_sched_mode(mode: int):
    L.check(L.lib().rt_set_sched_mode(mode))


# ---- host mirror hooks (include/rt_b200_host.h): the reference's Camera / projection / OBJ loading, host only
def host_camera(eye=(0.0, 0.0, 2.5), target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), lock_up=True, n_rotate=0, dx=0.0,
                dy=0.0):
    """Camera(eye, target, up) + setLockUp + n x rotate(dx, dy) -> (position[3], view[16], viewInv[16])."""
    a = [np.ascontiguousarray(x, np.float32) for x in (eye, target, up)]
    pos, view, vinv = np.zeros(3, np.float32), np.zeros(16, np.float32), np.zeros(16, np.float32)
    L.check(L.lib().rt_host_camera(_ptr(a[0]), _ptr(a[1]), _ptr(a[2]), int(lock_up), int(n_rotate), float(dx),
                                   float(dy), _ptr(pos), _ptr(view), _ptr(vinv)))
    return pos, view, vinv


def host_projection(W, H, fovy=45.0, z_near=0.01, z_far=100.0):
    """main.cpp:198-201 -> (proj[16], projInv[16])."""
    proj, inv = np.zeros(16, np.float32), np.zeros(16, np.float32)
    aspect = float(np.float32(W) / np.float32(H))
    L.check(L.lib().rt_host_projection(fovy, aspect, z_near, z_far, _ptr(proj), _ptr(inv)))
    return proj, inv


def host_load_obj(path: str, scale: bool = True):
    """LoadMeshFromObj (+ loadAndScale) -> (vpos4 [n,4] f32, indices [m] u32)."""
    pv, pi = C.POINTER(C.c_float)(), C.POINTER(C.c_uint32)()
    nv, ni = C.c_uint32(), C.c_uint32()
    if L.lib().rt_host_load_obj(path.encode(), int(scale), C.byref(pv), C.byref(nv), C.byref(pi), C.byref(ni)) != 0:
        raise L.RtError(f"cannot load {path}")
    try:
        v = np.ctypeslib.as_array(pv, shape=(nv.value, 4)).copy()
        i = np.ctypeslib.as_array(pi, shape=(ni.value,)).copy()
    finally:
        L.lib().rt_host_free(pv)
        L.lib().rt_host_free(pi)
    return v, i
