"""Generates tests/golden/small_meshes.npz from the reference itself (run HERE, where /root/reference exists).

    make -C oracle ref && python tests/golden/make_small_mesh_golden.py

Small meshes that are NOT a cube, rendered by the reference's unmodified BVHBuilder + Renderer::draw
(-ffp-contract=off build) from outside, grazing, inside and looking away, into cleared and pre-filled framebuffers,
with the plain Normal shading and with the viewer's default look (Lambert + shadows + reflections + ground plane).
Up to 8 triangles the reference's tree is ONE leaf (triangles_raytracing.cpp:175-186) whose triangles are tested with no
box and no range check (:292-331); beyond that the answer from inside / with pre-filled depth depends on the tree, so the
reference-built tree travels with the fixture (for rt_scene_create_mesh_bvh8) and the views are labelled.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_oracle as R  # noqa: E402

W, H = 48, 32


def _mesh(verts, tris):
    v = np.asarray(verts, np.float32)
    v4 = np.concatenate([v, np.ones((len(v), 1), np.float32)], 1)
    return v4, np.asarray(tris, np.uint32).reshape(-1)


def quad():  # 2 triangles, tilted, not planar
    return _mesh([(-0.6, -0.5, 0.1), (0.7, -0.4, -0.2), (0.6, 0.6, 0.3), (-0.5, 0.5, -0.1)], [(0, 1, 2), (0, 2, 3)])


def tetra():  # closed, 4 triangles
    return _mesh([(0.6, 0.6, 0.6), (-0.6, -0.6, 0.6), (-0.6, 0.6, -0.6), (0.6, -0.6, -0.6)],
                 [(0, 1, 2), (0, 3, 1), (0, 2, 3), (1, 3, 2)])


def fan():  # open fan, 5 triangles
    vs = [(0, 0.1, 0.2)] + [(0.7 * np.cos(a), 0.7 * np.sin(a), 0.15 * np.sin(3 * a)) for a in np.linspace(0, 4.2, 6)]
    return _mesh(vs, [(0, i, i + 1) for i in range(1, 6)])


def blob32():  # octahedron subdivided once, radii modulated: 32 triangles, closed
    v = [np.array(x, float) for x in [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]]
    t = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]
    nt, cache = [], {}

    def mid(a, b):
        k = (min(a, b), max(a, b))
        if k not in cache:
            m = (v[a] + v[b]) / 2
            v.append(m / np.linalg.norm(m))
            cache[k] = len(v) - 1
        return cache[k]

    for a, b, c in t:
        ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
        nt += [(a, ab, ca), (ab, b, bc), (ca, bc, c), (ab, bc, ca)]
    v = [x * (0.55 + 0.25 * np.sin(3 * x[0] + 2 * x[1]) * np.cos(2 * x[2])) for x in v]
    return _mesh(v, nt)


def spot32():  # the reference's spot.obj decimated to <= 32 triangles by vertex clustering on a coarse lattice
    d = np.load(os.path.join(HERE, "scenes", "spot.npz"))
    p = d["vpos4"][:, :3] / d["vpos4"][:, 3:4]
    tri = d["indices"].reshape(-1, 3)
    for cells in (2.2, 2.0, 1.8, 1.6, 1.4, 1.2):
        key = np.floor((p - p.min(0)) / (np.ptp(p, axis=0) + 1e-6) * cells).astype(np.int64)
        ids, inv = np.unique(key, axis=0, return_inverse=True)
        inv = inv.reshape(-1)
        rep = np.stack([p[inv == k].mean(0) for k in range(len(ids))])
        t = inv[tri]
        t = t[(t[:, 0] != t[:, 1]) & (t[:, 1] != t[:, 2]) & (t[:, 0] != t[:, 2])]
        canon = np.sort(t, axis=1)
        _, first = np.unique(canon, axis=0, return_index=True)
        t = t[np.sort(first)]
        if 9 <= len(t) <= 32:
            return _mesh(rep, t)
    raise RuntimeError("no lattice gave 9..32 triangles")


MESHES = dict(quad=quad, tetra=tetra, fan=fan, blob32=blob32, spot32=spot32)
# (label, eye radii, looks away from the mesh)
VIEWS = [("outside", (1.8, 2.5, 4.0), False), ("outside", (1.8, 2.5, 4.0), False), ("grazing", (1.0, 1.2), False),
         ("grazing", (1.0, 1.2), False), ("inside", (0.1, 0.3, 0.5), False), ("inside", (0.1, 0.3, 0.5), False),
         ("away", (1.0, 1.8, 2.5), True), ("inside_away", (0.1, 0.3), True)]


def main():
    o = R.RefOracle()
    pinv = o.projection(W, H)[1]
    out = {"projInv": pinv, "W": np.int32(W), "H": np.int32(H), "meshes": np.array(sorted(MESHES))}
    for name in sorted(MESHES):
        rng = np.random.default_rng(sum(name.encode()))
        v4, idx = MESHES[name]()
        ref = o.mesh(v4, idx)
        m = ref.export_mesh()
        out[f"{name}|vpos4"], out[f"{name}|indices"] = v4, idx
        out[f"{name}|indices_bvh"], out[f"{name}|nodes"] = m["indices"], m["nodes"]
        out[f"{name}|prim_of_permuted"] = m["prim_of_permuted"]
        plane_y = ref.model_min_y  # main.cpp:176-189
        out[f"{name}|plane_y"] = np.float32(plane_y)
        full = ref.with_plane(plane_y)
        labels = []
        for k, (label, radii, away) in enumerate(VIEWS):
            vv = rng.normal(size=3)
            vv /= np.linalg.norm(vv)
            r = float(rng.choice(radii))
            eye = tuple(float(x) for x in r * vv)
            target = (tuple(float(x) for x in 2.5 * max(r, 0.4) * vv + rng.uniform(-0.5, 0.5, 3)) if away
                      else tuple(float(x) for x in rng.uniform(-0.3, 0.3, 3)))
            cam = R.CamSpec(eye=eye, target=target, lock_up=bool(k % 2))
            pos, _view, vinv = o.camera(cam)
            t0 = rng.choice(np.array([np.inf, 0.005, 0.5, 1.2, 2.0, 2.6, 3.5], np.float32), size=(H, W)).astype(np.float32)
            c0 = ((np.arange(H * W, dtype=np.uint32) % 251) * np.uint32(0x01020301) + np.uint32(k)).reshape(H, W)
            key = f"{name}|{k}"
            labels.append(label)
            out[key + "|rayPos"], out[key + "|viewInv"] = pos, vinv
            out[key + "|t0"], out[key + "|c0"] = t0, c0
            out[key + "|color"], out[key + "|t"], _ = ref.render(cam, W, H, proj_inv=pinv)
            out[key + "|color_pre"], out[key + "|t_pre"], _ = ref.render(cam, W, H, proj_inv=pinv, color=c0, t=t0)
            out[key + "|color_full"], out[key + "|t_full"], _ = full.render(cam, W, H, proj_inv=pinv, mode=R.MODE_LAMBERT,
                                                                           shadows=True, reflections=True)
            out[key + "|color_full_pre"], out[key + "|t_full_pre"], _ = full.render(
                cam, W, H, proj_inv=pinv, mode=R.MODE_LAMBERT, shadows=True, reflections=True, color=c0, t=t0)
        out[f"{name}|views"] = np.array(labels)
        print(f"{name}: {idx.size // 3} triangles, reference tree of {m['nodes'].shape[0]} nodes, plane_y {plane_y:.4f}")
    np.savez_compressed(os.path.join(HERE, "small_meshes.npz"), **out)


if __name__ == "__main__":
    main()
