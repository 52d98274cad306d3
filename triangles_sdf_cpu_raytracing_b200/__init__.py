"""B200-native primary-ray hot path of iMacsimus/Triangles-SDF-CPU-RayTracing.

The product is librt_b200.so (CUDA, sm_100a) behind include/rt_b200.h; this package is its Python face.
"""
from ._lib import (GRID_AUTO, GRID_BRICKS, GRID_LINEAR, LIB_PATH, MATH_EXACT, MATH_FAST, MESH_AUTO, MESH_BRUTE,
                   MESH_BVH8, MESH_LBVH_GPU, PRIM_MISS, PRIM_PLANE, SCHED_AUTO, SCHED_COMPACT, SCHED_PERSISTENT, SCHED_TILES, SHADE_COLOR,
                   SHADE_LAMBERT, SHADE_NORMAL, RtCounters, RtError, RtFrame, lib)
from .api import (PinnedFrameBuffer, host_camera, host_load_obj, host_projection, Scene, SDFGrid, SDFOctree, TriangleMesh, TriangleMeshFromReferenceBVH, load_grid_file,
                  init, load_octree_file, make_frame, set_math_mode, set_sched_mode)
