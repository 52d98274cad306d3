"""ctypes binding of librt_b200.so (include/rt_b200.h). There is no fallback: a missing library is an error."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_B200_LIB selects another build of the same library (kernel experiments, tools/build_variants.sh)
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")

RT_OK = 0
SHADE_NORMAL, SHADE_LAMBERT, SHADE_COLOR = 0, 1, 2
MATH_EXACT, MATH_FAST = 0, 1
SCHED_AUTO, SCHED_TILES, SCHED_PERSISTENT, SCHED_COMPACT = 0, 1, 2, 3
GRID_AUTO, GRID_LINEAR, GRID_BRICKS = 0, 1, 2
MESH_AUTO, MESH_BVH8, MESH_BRUTE, MESH_LBVH_GPU = 0, 1, 2, 3
SCENE_GRID, SCENE_OCTREE, SCENE_MESH = 0, 1, 2
PRIM_MISS = 0xFFFFFFFF
PRIM_PLANE = 0xFFFFFFFE


class RtFrame(C.Structure):
    _fields_ = [
        ("ray_pos", C.c_float * 3), ("view_inv", C.c_float * 16), ("proj_inv", C.c_float * 16),
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("band_rows", C.c_uint32), ("band_count", C.c_uint32), ("band_index", C.c_uint32),
        ("shading_mode", C.c_int32), ("light_pos", C.c_float * 3),
        ("enable_shadows", C.c_int32), ("enable_reflections", C.c_int32),
        ("has_plane", C.c_int32), ("plane_y", C.c_float),
        ("t_near", C.c_float), ("t_far", C.c_float), ("clear", C.c_int32),
    ]


class RtCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "hits", "grid_samples", "oct_inner", "oct_leaf",
                                          "oct_leaf_samples", "bvh_inner", "bvh_leaf", "bvh_tris")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


# every symbol include/rt_b200.h declares: name -> (restype, argtypes)
_vp, _u32, _i, _f = C.c_void_p, C.c_uint32, C.c_int, C.c_float
_pp = C.POINTER(C.c_void_p)
SIGNATURES = {
    "rt_last_error": (C.c_char_p, []),
    "rt_version": (C.c_char_p, []),
    "rt_device_count": (_i, [C.POINTER(_i)]),
    "rt_set_device": (_i, [_i]),
    "rt_init": (_i, [_vp, _i]),
    "rt_device_list": (_i, [_vp, _i]),
    "rt_scene_replicas": (_i, [_vp, _vp, _i]),
    "rt_scene_create_grid": (_i, [_vp, _u32, _u32, _u32, _i, _pp]),
    "rt_scene_create_octree": (_i, [_vp, _u32, _pp]),
    "rt_scene_create_mesh": (_i, [_vp, _u32, _vp, _u32, _i, _pp]),
    "rt_scene_create_mesh_bvh8": (_i, [_vp, _u32, _vp, _u32, _vp, _u32, _u32, _vp, _pp]),
    "rt_scene_build_ms": (_i, [_vp, C.POINTER(_f)]),
    "rt_scene_mesh_stats": (_i, [_vp, C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_f)]),
    "rt_scene_destroy": (None, [_vp]),
    "rt_scene_kind_of": (_i, [_vp]),
    "rt_scene_device": (_i, [_vp]),
    "rt_scene_info": (_i, [_vp, C.POINTER(C.c_uint64), C.POINTER(_f), C.POINTER(_f)]),
    "rt_scene_transfer_bytes": (_i, [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "rt_set_math_mode": (_i, [_i]),
    "rt_set_sched_mode": (_i, [_i]),
    "rt_render_device": (_i, [_vp, C.POINTER(RtFrame), _vp, _vp, _vp, _vp, _vp, _vp]),
    "rt_render_host": (_i, [_vp, C.POINTER(RtFrame), _vp, _vp, _vp, C.POINTER(_f), C.POINTER(_f),
                            C.POINTER(RtCounters)]),
    "rt_intersect_host": (_i, [_vp, _vp, _vp, _u32, _f, _f, _i, _f, _vp, _vp, _vp, _vp]),
    "rt_intersect_host_static": (_i, [_vp, _vp, _vp, _u32, _f, _f, _i, _f, _vp, _vp, _vp, _vp]),
    "rt_device_malloc": (_i, [C.c_uint64, _pp]),
    "rt_device_free": (_i, [_vp]),
    "rt_ipc_export": (_i, [_vp, C.c_char_p]),
    "rt_ipc_open": (_i, [C.c_char_p, _pp]),
    "rt_ipc_close": (_i, [_vp]),
    "rt_staged_launch_count": (C.c_uint64, []),
    "rt_pinned_alloc": (_i, [C.c_uint64, _pp]),
    "rt_pinned_free": (_i, [_vp]),
    "rt_pinned_register": (_i, [_vp, C.c_uint64]),
    "rt_pinned_unregister": (_i, [_vp]),
    "rt_memcpy_d2h": (_i, [_vp, _vp, C.c_uint64, _vp]),
    "rt_memcpy_h2d": (_i, [_vp, _vp, C.c_uint64, _vp]),
    # include/rt_b200_host.h
    "rt_host_camera": (_i, [_vp, _vp, _vp, _i, _i, _f, _f, _vp, _vp, _vp]),
    "rt_host_projection": (_i, [_f, _f, _f, _f, _vp, _vp]),
    "rt_host_load_obj": (_i, [C.c_char_p, _i, C.POINTER(C.POINTER(_f)), C.POINTER(_u32),
                              C.POINTER(C.POINTER(_u32)), C.POINTER(_u32)]),
    "rt_host_free": (None, [_vp]),
}

_lib = None


class RtError(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Loads the CUDA library. Raises if it has not been built (`make` or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(f"{LIB_PATH} is missing: build it with `make` at the repository root "
                          "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int):
    if rc != RT_OK:
        raise RtError(f"rt_b200 error {rc}: {lib().rt_last_error().decode()}")
