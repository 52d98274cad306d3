"""Shared helpers of the test-suite: golden fixtures, oracle scene construction, comparisons."""
from __future__ import annotations

import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
MESHES = ("cube", "stanford-bunny", "spot", "as1-oc-214")


class Golden:
    def __init__(self):
        self.frames = np.load(os.path.join(GOLDEN, "frames.npz"))
        self.kats = np.load(os.path.join(GOLDEN, "kats.npz"))
        self.cases = [str(c) for c in self.frames["cases"]]
        self._scenes = {}

    def scene_arrays(self, name):
        """Fixture copy of a shipped resource, or a deterministic stand-in derived from one (standins.py):
        example_grid_x4 / _x8 (257^3 / 513^3 refinements), octree_d8 / octree_d9 (octrees of those grids),
        stanford-bunny_s1 / _s2 (x4 / x16 midpoint subdivisions)."""
        if name not in self._scenes:
            from triangles_sdf_cpu_raytracing_b200 import standins as S
            if name.startswith("example_grid_x"):
                a = dict(values=S.refine_grid(self.scene_arrays("example_grid")["values"], int(name.rsplit("x", 1)[1])))
            elif name.startswith("octree_d"):
                d = int(name.rsplit("d", 1)[1])
                grid = S.refine_grid(self.scene_arrays("example_grid")["values"], (1 << d) // 64)
                a = dict(nodes36=S.octree_from_grid(grid, d))
            elif name.startswith("stanford-bunny_s"):
                m = self.scene_arrays("stanford-bunny")
                v, i = S.subdivide_mesh(m["vpos4"], m["indices"], int(name.rsplit("s", 1)[1]))
                a = dict(vpos4=v, indices=i)
            else:
                a = dict(np.load(os.path.join(GOLDEN, "scenes", f"{name}.npz")))
            self._scenes[name] = a
        return self._scenes[name]

    def case(self, key):
        scene, cam, wh = key.split("|")
        W, H = (int(x) for x in wh.split("x"))
        g = lambda n: self.frames[f"{key}|{n}"] if f"{key}|{n}" in self.frames else None  # noqa: E731
        return dict(scene=scene, cam=cam, W=W, H=H, rayPos=g("rayPos"), viewInv=g("viewInv"), projInv=g("projInv"),
                    color_normal=g("color_normal"), color_lambert=g("color_lambert"), color_color=g("color_color"),
                    t=g("t"), prim=g("prim"), color_prefilled=g("color_prefilled"), t_prefilled=g("t_prefilled"),
                    color_full=g("color_full"), t_full=g("t_full"), plane_y=g("plane_y"), color_shadow=g("color_shadow"))


def port_scene(golden: Golden, name: str):
    """The plain-C oracle's scene for a golden scene name (meshes: brute force, tree-independent)."""
    from oracle import port_oracle as P
    a = golden.scene_arrays(name)
    if "values" in a:
        return P.Scene.grid(a["values"])
    if "nodes36" in a:
        return P.Scene.octree(a["nodes36"])
    return P.Scene.mesh_brute(a["vpos4"], a["indices"])


def cuda_scene(golden: Golden, name: str, **kw):
    import triangles_sdf_cpu_raytracing_b200 as rt
    a = golden.scene_arrays(name)
    if "values" in a:
        return rt.SDFGrid(a["values"], **kw)
    if "nodes36" in a:
        return rt.SDFOctree(a["nodes36"])
    return rt.TriangleMesh(a["vpos4"], a["indices"], **kw)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_bits(a, b):
    return np.array_equal(bits(a), bits(b))


def rgb_max_diff(c0, c1):
    """max per-channel absolute difference of two packed RGBA images (in 1/255 units)."""
    a = np.ascontiguousarray(c0, np.uint32).view(np.uint8).astype(np.int32)
    b = np.ascontiguousarray(c1, np.uint32).view(np.uint8).astype(np.int32)
    return int(np.abs(a - b).max()) if a.size else 0


def compare_frames(color, t, ref_color, ref_t, depth_rtol=1e-4):
    """The north-star tolerances: hit mask exact bar counted pixels, depth 1e-4 relative, RGB within 1/255.
    -> dict(mask_mismatch, depth_outliers, rgb_outliers, n_hit)."""
    hit, rhit = np.isfinite(t), np.isfinite(ref_t)
    both = hit & rhit
    rel = np.abs(t[both].astype(np.float64) - ref_t[both]) / np.abs(ref_t[both].astype(np.float64))
    a = np.ascontiguousarray(color, np.uint32).view(np.uint8).reshape(color.shape + (4,)).astype(np.int32)
    b = np.ascontiguousarray(ref_color, np.uint32).view(np.uint8).reshape(color.shape + (4,)).astype(np.int32)
    rgb = np.abs(a - b).max(axis=-1)
    return dict(mask_mismatch=int((hit != rhit).sum()), depth_outliers=int((rel > depth_rtol).sum()),
                rgb_outliers=int((rgb[both] > 1).sum()), n_hit=int(rhit.sum()),
                max_rel_depth=float(rel.max()) if rel.size else 0.0)
