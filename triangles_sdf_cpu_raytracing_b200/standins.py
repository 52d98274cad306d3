"""Deterministic stand-ins for the three benchmark resources the reference tree does not ship
(`example_grid_large.grid`, `example_octree_large.octree`, `MotorcycleCylinderHead.obj`).

Nothing here uses a random number generator: every stand-in is a pure function of a resource the reference does
ship, written in the reference's own file formats (Appendix F of SURVEY.md), so the reference's loaders
(grid_raytracing.cpp:127-134, octree_raytracing.cpp:8-16, mesh.cpp:178-286) would read them unchanged.

  refine_grid      example_grid.grid (65^3) -> 257^3 / 513^3 by exact trilinear refinement: the refined grid
                   samples the SAME piecewise-trilinear function, so the picture is the small grid's picture
                   (up to rounding) while the data no longer fits L1 (257^3 = 68 MB: L2) or L2 (513^3 = 540 MB: HBM).
  octree_from_grid a (2^d + 1)^3 grid -> an SDF octree of depth d following the conventions observed in the shipped
                   sdf_5/sdf_6.octree files (octree_raytracing.hpp:8-18): inner nodes carry all-zero values, the 8
                   children of a node are one contiguous block at childrenOffset + ((x<<2)|(y<<1)|z), leaves that
                   hold no surface are all-1000.0 (outside) or all-0.0 (inside), surface leaves hold the 8 corner
                   samples in (x<<2)+(y<<1)+z order and sit at the maximum depth.
  subdivide_mesh   midpoint (1 -> 4) subdivision of a triangle mesh, `levels` times; triangle k of the input becomes
                   triangles 4k .. 4k+3 of the output, so prim ids stay meaningful.
"""
from __future__ import annotations

import numpy as np


# ------------------------------------------------------------------------------------------------- dense grid
def _refine_axis(a: np.ndarray, axis: int, factor: int) -> np.ndarray:
    """Linear refinement of `a` (float64) along one axis: n samples -> (n-1)*factor + 1."""
    a = np.moveaxis(a, axis, 0)
    n = a.shape[0]
    out = np.empty(((n - 1) * factor + 1,) + a.shape[1:], np.float64)
    for k in range(factor):
        w = k / factor  # exact in binary for the power-of-two factors used here
        out[k:-1:factor] = a[:-1] * (1.0 - w) + a[1:] * w if k else a[:-1]
    out[-1] = a[-1]
    return np.moveaxis(out, 0, axis)


def refine_grid(values: np.ndarray, factor: int) -> np.ndarray:
    """values[x, y, z] float32 with s samples per axis -> ((s-1)*factor + 1)^3 float32, the trilinear interpolant of
    `values` (grid_raytracing.cpp:7-62) evaluated in float64 at the refined sample positions and rounded once."""
    v = np.ascontiguousarray(values, np.float32)
    if factor == 1:
        return v.copy()
    yz = _refine_axis(_refine_axis(v.astype(np.float64), 2, factor), 1, factor)  # [s, S, S]
    s = v.shape[0]
    S = (s - 1) * factor + 1
    out = np.empty((S,) + yz.shape[1:], np.float32)
    for i in range(s - 1):  # slab by slab: keeps the float64 working set small for the 513^3 case
        for k in range(factor):
            w = k / factor
            out[i * factor + k] = (yz[i] * (1.0 - w) + yz[i + 1] * w) if k else yz[i]
    out[-1] = yz[-1]
    return out


def grid_file_bytes(values: np.ndarray) -> bytes:
    """The .grid image loadSDFGrid reads (grid_raytracing.cpp:127-134)."""
    v = np.ascontiguousarray(values, np.float32)
    return np.asarray(v.shape, np.uint32).tobytes() + v.tobytes()


# ------------------------------------------------------------------------------------------------- octree
def octree_from_grid(values: np.ndarray, depth: int | None = None) -> np.ndarray:
    """-> uint8 [n, 36]: the raw SDFOctreeNode records of an octree whose depth-`depth` leaves are the cells of the
    (2^depth + 1)^3 grid `values`.  A node is split while its block of cells holds a sign change of the samples."""
    v = np.ascontiguousarray(values, np.float32)
    n = v.shape[0] - 1
    if depth is None:
        depth = int(round(np.log2(n)))
    if v.shape != (n + 1,) * 3 or (1 << depth) != n:
        raise ValueError("octree_from_grid needs a (2^depth + 1)^3 grid")
    # per-cell min / max of the 8 corners, then a min/max pyramid over 2x2x2 blocks
    c = [v[dx:n + dx, dy:n + dy, dz:n + dz] for dx in (0, 1) for dy in (0, 1) for dz in (0, 1)]  # (x<<2)+(y<<1)+z
    lo = [np.minimum.reduce(c)]
    hi = [np.maximum.reduce(c)]
    for _ in range(depth):
        a, b = lo[-1], hi[-1]
        m = a.shape[0] // 2
        lo.append(a.reshape(m, 2, m, 2, m, 2).min(axis=(1, 3, 5)))
        hi.append(b.reshape(m, 2, m, 2, m, 2).max(axis=(1, 3, 5)))
    lo.reverse()
    hi.reverse()  # index = level: level 0 is the root (1 block), level `depth` the cells

    vals_per_level, off_per_level = [], []
    coords = np.zeros((1, 3), np.int64)  # nodes of the current level, in node-id order
    first_id = 0                         # id of the first node of the current level
    for level in range(depth + 1):
        ix, iy, iz = coords[:, 0], coords[:, 1], coords[:, 2]
        mn, mx = lo[level][ix, iy, iz], hi[level][ix, iy, iz]
        surface = (mn < 0.0) & (mx > 0.0)
        count = coords.shape[0]
        vals = np.zeros((count, 8), np.float32)
        off = np.zeros(count, np.uint32)
        if level == depth:
            for k in range(8):
                vals[:, k] = c[k][ix, iy, iz]
            vals[~surface & (mn >= 0.0)] = 1000.0  # empty outside
            vals[~surface & (mn < 0.0)] = 0.0      # empty inside
        else:
            inner = surface
            vals[~inner & (mn >= 0.0)] = 1000.0
            vals[~inner & (mn < 0.0)] = 0.0
            n_inner = int(inner.sum())
            next_first = first_id + count
            off[inner] = (next_first + 8 * np.arange(n_inner)).astype(np.uint32)
            kid = np.arange(8)
            d = np.stack([(kid >> 2) & 1, (kid >> 1) & 1, kid & 1], axis=1)  # child id -> (x, y, z) bit
            coords = (2 * coords[inner][:, None, :] + d[None, :, :]).reshape(-1, 3)
            first_id = next_first
        vals_per_level.append(vals)
        off_per_level.append(off)
        if level < depth and coords.shape[0] == 0:
            break
    vals = np.concatenate(vals_per_level)
    off = np.concatenate(off_per_level)
    rec = np.zeros((vals.shape[0], 36), np.uint8)
    rec[:, :32] = vals.view(np.uint8).reshape(-1, 32)
    rec[:, 32:] = off.view(np.uint8).reshape(-1, 4)
    return rec


def octree_file_bytes(nodes36: np.ndarray) -> bytes:
    """The .octree image loadSDFOctree reads (octree_raytracing.cpp:8-16)."""
    r = np.ascontiguousarray(nodes36, np.uint8).reshape(-1, 36)
    return np.asarray([r.shape[0]], np.uint32).tobytes() + r.tobytes()


# ------------------------------------------------------------------------------------------------- mesh
def subdivide_mesh(vpos4: np.ndarray, indices: np.ndarray, levels: int = 1):
    """Midpoint subdivision. (vpos4 [n,4] f32, indices [3m] u32) -> the same with 4^levels x the triangles.
    Corner triangles first, centre triangle last: triangle k -> 4k (a,ab,ca), 4k+1 (ab,b,bc), 4k+2 (ca,bc,c),
    4k+3 (ab,bc,ca); winding preserved; midpoints are shared between the two triangles of an edge."""
    v = np.ascontiguousarray(vpos4, np.float32).reshape(-1, 4)
    idx = np.ascontiguousarray(indices, np.uint32).reshape(-1, 3).astype(np.int64)
    for _ in range(levels):
        nv = v.shape[0]
        e = np.concatenate([idx[:, [0, 1]], idx[:, [1, 2]], idx[:, [2, 0]]])  # ab, bc, ca
        key = np.minimum(e[:, 0], e[:, 1]) * nv + np.maximum(e[:, 0], e[:, 1])
        uniq, inv = np.unique(key, return_inverse=True)
        lo, hi = uniq // nv, uniq % nv
        mid = ((v[lo] + v[hi]) * np.float32(0.5)).astype(np.float32)
        mid[:, 3] = 1.0
        m = idx.shape[0]
        ab, bc, ca = nv + inv[:m], nv + inv[m:2 * m], nv + inv[2 * m:]
        a, b, c = idx[:, 0], idx[:, 1], idx[:, 2]
        idx = np.stack([np.stack([a, ab, ca], 1), np.stack([ab, b, bc], 1), np.stack([ca, bc, c], 1),
                        np.stack([ab, bc, ca], 1)], axis=1).reshape(-1, 3)
        v = np.concatenate([v, mid])
    return v, idx.reshape(-1).astype(np.uint32)
