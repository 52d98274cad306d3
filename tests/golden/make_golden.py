"""Generates tests/golden/ from the reference itself (run HERE, where /root/reference exists).

    make -C oracle ref && python tests/golden/make_golden.py

Outputs (all committed; the GPU box has no /root/reference):
  scenes/<name>.npz   the reference's resource files converted to arrays exactly as the reference's own
                      loaders see them (loadSDFGrid grid_raytracing.cpp:127-134, loadSDFOctree
                      octree_raytracing.cpp:8-16, LoadMeshFromObj mesh.cpp:178-286 + loadAndScale
                      main.cpp:326-343).  Meshes are stored in LoadMeshFromObj order (= prim ids).
  frames.npz          full frames rendered by the reference's unmodified Renderer::draw
                      (-ffp-contract=off build): colour, depth and, for meshes, instrumented prim ids, for a
                      set of (scene, camera, mode) cases, together with the camera matrices that produced
                      them (computed by the reference's Camera / perspectiveMatrix / inverse4x4).
  kats.npz            known-answer vectors for the primitives (slab test, BBox3f::Intersection, sort8,
                      Moeller-Trumbore, trilinear sample, normals, eye rays, colour pack) on seeded inputs
                      plus hand-picked edge cases (axis-parallel rays, origins on slabs, integer coords).
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_oracle as R  # noqa: E402

RES = "/root/reference/resources"
ORBIT_DX = float(np.float32(2.0 * np.pi / 64.0 / 0.01))  # 64-frame orbit step


def cameras():
    return {
        "default": R.CamSpec(),
        "orbit5": R.CamSpec(n_rotate=5, dx=ORBIT_DX, dy=0.0),
        "orbit23tilt": R.CamSpec(n_rotate=23, dx=ORBIT_DX, dy=1.5),
        "plusx": R.CamSpec(eye=(2.5, 0.0, 0.0)),  # numpad side view: axis-parallel centre rays
        "close": R.CamSpec(eye=(0.3, 0.2, 1.2)),
    }


def save_scenes(o: R.RefOracle):
    out = os.path.join(HERE, "scenes")
    os.makedirs(out, exist_ok=True)
    raw = np.fromfile(f"{RES}/example_grid.grid", dtype=np.uint8)
    size = raw[:12].view(np.uint32)
    vals = raw[12:].view(np.float32).reshape(tuple(int(x) for x in size))
    np.savez_compressed(f"{out}/example_grid.npz", values=vals)
    for name in ("sdf_5", "sdf_6"):
        raw = np.fromfile(f"{RES}/{name}.octree", dtype=np.uint8)
        n = int(raw[:4].view(np.uint32)[0])
        np.savez_compressed(f"{out}/{name}.npz", nodes36=raw[4:4 + 36 * n].reshape(n, 36))
    for name in ("cube", "stanford-bunny", "spot", "as1-oc-214"):
        s = o.mesh_obj(f"{RES}/{name}.obj", scale=True)
        m = s.export_mesh()
        perm = m["prim_of_permuted"]
        idx_bvh = m["indices"].reshape(-1, 3)
        idx = np.zeros_like(idx_bvh)
        idx[perm] = idx_bvh  # back to LoadMeshFromObj order
        np.savez_compressed(f"{out}/{name}.npz", vpos4=m["vpos4"], indices=idx.reshape(-1))
        if name == "cube":  # the reference-built BVH8 (204-byte nodes), for the import path's tests
            np.savez_compressed(f"{out}/cube_refbvh.npz", vpos4=m["vpos4"], indices_bvh=m["indices"],
                                nodes=m["nodes"], prim_of_permuted=perm)


def load_scene(o: R.RefOracle, name: str):
    d = np.load(os.path.join(HERE, "scenes", f"{name}.npz"))
    if "values" in d:
        return o.grid(d["values"])
    if "nodes36" in d:
        return o.octree(d["nodes36"])
    return o.mesh(d["vpos4"], d["indices"], scale=False)


def save_frames(o: R.RefOracle):
    cams = cameras()
    cases = [  # (scene, camera, W, H)
        ("cube", "default", 128, 128), ("cube", "orbit23tilt", 96, 64), ("cube", "plusx", 64, 64),
        ("example_grid", "default", 160, 90), ("example_grid", "orbit5", 96, 64),
        ("example_grid", "plusx", 65, 65), ("example_grid", "close", 96, 64),
        ("sdf_6", "default", 160, 90), ("sdf_6", "orbit23tilt", 96, 64), ("sdf_6", "plusx", 65, 65),
        ("sdf_5", "default", 96, 64), ("sdf_5", "close", 96, 64),
        ("stanford-bunny", "default", 160, 90), ("stanford-bunny", "orbit5", 96, 64),
        ("spot", "orbit23tilt", 96, 64), ("as1-oc-214", "default", 96, 64),
    ]
    out = {}
    names = []
    scenes = {}
    for scene, cam, W, H in cases:
        if scene not in scenes:
            scenes[scene] = load_scene(o, scene)
        s = scenes[scene]
        key = f"{scene}|{cam}|{W}x{H}"
        names.append(key)
        pos, _view, view_inv = o.camera(cams[cam])
        proj_inv = o.projection(W, H)[1]
        out[key + "|rayPos"] = pos
        out[key + "|viewInv"] = view_inv
        out[key + "|projInv"] = proj_inv
        col, t, _ = s.render(cams[cam], W, H, mode=R.MODE_NORMAL)
        out[key + "|color_normal"] = col
        out[key + "|t"] = t
        col, t2, _ = s.render(cams[cam], W, H, mode=R.MODE_LAMBERT, shadows=False, reflections=False)
        assert np.array_equal(t, t2)
        out[key + "|color_lambert"] = col
        col, _, _ = s.render(cams[cam], W, H, mode=R.MODE_COLOR)
        out[key + "|color_color"] = col
        if s.kind == R.KIND_MESH:
            tp, _n, prim = s.render_prim(cams[cam], W, H)
            assert np.array_equal(tp, t)
            out[key + "|prim"] = prim
        # the viewer's default look: Lambert + shadow rays + reflections + checkerboard ground plane
        if cam in ("default", "orbit5", "close"):
            plane_y = s.model_min_y if s.kind == R.KIND_MESH else -1.0  # main.cpp:176-189
            full = s.with_plane(plane_y)
            col, tf, _ = full.render(cams[cam], W, H, mode=R.MODE_LAMBERT, shadows=True, reflections=True)
            out[key + "|color_full"] = col
            out[key + "|t_full"] = tf
            out[key + "|plane_y"] = np.float32(plane_y)
            col, _, _ = s.render(cams[cam], W, H, mode=R.MODE_LAMBERT, shadows=True, reflections=False)
            out[key + "|color_shadow"] = col
        # in/out depth buffer: a pre-filled depth plane at t=2.2 hides farther hits
        t0 = np.full((H, W), 2.2, np.float32)
        c0 = np.full((H, W), 0x11223344, np.uint32)
        col, t3, _ = s.render(cams[cam], W, H, mode=R.MODE_NORMAL, color=c0, t=t0)
        out[key + "|color_prefilled"] = col
        out[key + "|t_prefilled"] = t3
    out["cases"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "frames.npz"), **out)


def save_kats(o: R.RefOracle):
    rng = np.random.default_rng(0)
    out = {}
    # ---- slab test on 8 boxes (ray_pack.ispc:241-273)
    N = 64
    lo = rng.uniform(-1, 0.5, (N, 3, 8)).astype(np.float32)
    hi = lo + rng.uniform(0.01, 1.0, (N, 3, 8)).astype(np.float32)
    boxes = np.concatenate([lo, hi], axis=1)  # [N, 6, 8]
    org = rng.uniform(-2, 2, (N, 3)).astype(np.float32)
    d = rng.normal(size=(N, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:8, 0] = 0.0  # axis-parallel rays: invDir = +inf
    d[8:12, 1] = -0.0
    org[:4, 0] = boxes[:4, 0, 0]  # origin exactly on a slab plane with zero direction: 0*inf = NaN
    with np.errstate(divide="ignore"):
        inv = (np.float32(1.0) / d).astype(np.float32)
    res = np.stack([o.intersect_box_8(boxes[i], org[i], inv[i], 0.01, 100.0) for i in range(N)])
    out.update(slab_boxes=boxes, slab_org=org, slab_inv=inv, slab_t=res)
    # ---- BBox3f::Intersection (LiteMath.h:2005-2067)
    t1 = np.zeros(N, np.float32)
    t2 = np.zeros(N, np.float32)
    face = np.zeros(N, np.int32)
    for i in range(N):
        t1[i], t2[i], face[i] = o.bbox_intersection(boxes[i, :3, 0], boxes[i, 3:, 0], org[i], inv[i], 0.01, 100.0)
    out.update(bbox_t1=t1, bbox_t2=t2, bbox_face=face)
    # ---- sort8 (raytracing.hpp:188-213), with ties and misses
    ts = rng.uniform(0, 4, (N, 8)).astype(np.float32)
    ts[rng.uniform(size=(N, 8)) < 0.3] = -1.0
    ts[rng.uniform(size=(N, 8)) < 0.2] = 0.01
    st = np.zeros_like(ts)
    sc = np.zeros((N, 8), np.int64)
    for i in range(N):
        st[i], sc[i] = o.sort8(ts[i])
    out.update(sort_in=ts, sort_t=st, sort_idx=sc)
    # ---- 1 ray x 8 triangles (ray_pack.ispc:289-310)
    tris = rng.uniform(-1, 1, (N, 8, 3, 3)).astype(np.float32)
    tris[0, 0, 2] = tris[0, 0, 1]  # degenerate triangle: NaN normal, rejected by |det| < 1e-8
    tro = rng.uniform(-2, 2, (N, 3)).astype(np.float32)
    trd = -tro + rng.normal(scale=0.3, size=(N, 3)).astype(np.float32)
    trd /= np.linalg.norm(trd, axis=1, keepdims=True)
    trd = trd.astype(np.float32)
    th = np.zeros((N, 8), np.uint32)
    tt = np.zeros((N, 8), np.float32)
    tn = np.zeros((N, 8, 3), np.float32)
    for i in range(N):
        th[i], tt[i], tn[i] = o.intersect_1_ray_8_triangles(tris[i], tro[i], trd[i])
    out.update(tri_v=tris, tri_o=tro, tri_d=trd, tri_hit=th, tri_t=tt, tri_n=tn)
    # ---- grid sample + normal (grid_raytracing.cpp:7-89)
    g = load_scene(o, "example_grid")
    P = rng.uniform(-1, 1, (256, 3)).astype(np.float32)
    P[:27] = np.array([[x, y, z] for x in (-1, 0, 1) for y in (-1, 0, 1) for z in (-1, 0, 1)], np.float32)
    P[27:40] = np.float32(-1.0) + np.float32(2.0) * rng.integers(0, 65, (13, 3)).astype(np.float32) / np.float32(64)
    P[40:48, 0] = 1.0
    out.update(grid_p=P, grid_sdf=np.array([g.grid_sdf(p) for p in P], np.float32),
               grid_normal=np.stack([g.grid_normal(p) for p in P]))
    # ---- single rays through every scene type (IScene::intersect)
    for name in ("example_grid", "sdf_6", "stanford-bunny", "cube"):
        s = load_scene(o, name)
        M = 192
        ro = rng.normal(size=(M, 3)).astype(np.float32)
        ro = (ro / np.linalg.norm(ro, axis=1, keepdims=True) * rng.uniform(1.8, 3.0, (M, 1))).astype(np.float32)
        rd = -ro + rng.normal(scale=0.35, size=(M, 3)).astype(np.float32)
        rd = (rd / np.linalg.norm(rd, axis=1, keepdims=True)).astype(np.float32)
        ro[:6] = np.array([[2.5, 0, 0], [-2.5, 0, 0], [0, 2.5, 0], [0, -2.5, 0], [0, 0, 2.5], [0, 0, -2.5]], np.float32)
        rd[:6] = -ro[:6] / np.float32(2.5)  # exactly axis-parallel (two zero components)
        hit = np.zeros(M, np.uint8)
        t = np.zeros(M, np.float32)
        n = np.zeros((M, 3), np.float32)
        for i in range(M):
            h, t[i], n[i] = s.intersect(ro[i], rd[i])
            hit[i] = h
        out.update({f"ray_{name}_o": ro, f"ray_{name}_d": rd, f"ray_{name}_hit": hit, f"ray_{name}_t": t,
                    f"ray_{name}_n": n})
    # ---- eye rays + colour pack
    pos, _v, vinv = o.camera(cameras()["orbit23tilt"])
    pinv = o.projection(96, 64)[1]
    xy = np.array([[0, 0], [95, 63], [47, 31], [48, 32], [0, 63], [95, 0], [13, 57]], np.uint32)
    out.update(eye_viewInv=vinv, eye_projInv=pinv, eye_xy=xy,
               eye_dir=np.stack([o.eye_ray(int(x), int(y), 96, 64, pinv, vinv) for x, y in xy]))
    cols = rng.uniform(0, 1, (64, 4)).astype(np.float32)
    cols[0] = (1, 1, 1, 1)
    cols[1] = (0, 0, 0, 0)
    cols[2] = (0.5, 0.25, 0.999999, 1)
    out.update(pack_in=cols, pack_out=np.array([o.color_pack(c) for c in cols], np.uint32))
    np.savez_compressed(os.path.join(HERE, "kats.npz"), **out)


def main():
    o = R.RefOracle(contract=False)
    save_scenes(o)
    save_frames(o)
    save_kats(o)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
