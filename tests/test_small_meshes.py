"""Small meshes that are NOT a cube against frames the reference itself rendered (tests/golden/small_meshes.npz, made by
tests/golden/make_small_mesh_golden.py from oracle/_ref): a 2-triangle quad, a tetrahedron, an open 5-triangle fan, a
32-triangle blob and a 23-triangle decimated spot.obj; eye outside, grazing, inside, looking away; cleared and
pre-filled framebuffers; Normal shading and the viewer's default look (plane + shadows + reflections).

What must hold (triangles_raytracing.cpp:155-186, 266-335):
  * <= 8 triangles: the reference's tree is one leaf, tested without box or range check -> RT_MESH_AUTO (the brute-force
    kernel in BRUTE_ROOT_LEAF mode) and the oracle's every-triangle loop equal the reference in EVERY view, bit for bit;
  * any size, the reference's own tree imported (rt_scene_create_mesh_bvh8): equal in EVERY view, bit for bit;
  * 9+ triangles, our own tree (RT_MESH_AUTO): equal wherever the reference's answer does not depend on its tree
    (eye outside, cleared framebuffer, primary rays); the other views are tree-dependent in the reference itself and are counted,
    not asserted (profiles/r2_small_mesh_tree_dependence.txt).
CPU half: the oracle and the kernels' source compiled for the host (tests/host_emul).  GPU half: through the C ABI."""
from __future__ import annotations

import os

import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from oracle import port_oracle as P
from tests import host_emul as E
from tests import util

SM = np.load(os.path.join(util.GOLDEN, "small_meshes.npz"))
W, H = int(SM["W"]), int(SM["H"])
PINV = SM["projInv"]
MESHES = [str(m) for m in SM["meshes"]]
ROOT_LEAF = [m for m in MESHES if SM[f"{m}|indices"].size // 3 <= 8]
TREE = [m for m in MESHES if m not in ROOT_LEAF]
# (golden colour, golden depth, pre-filled buffers?, default look?)
RESULTS = [("color", "t", False, False), ("color_pre", "t_pre", True, False), ("color_full", "t_full", False, True),
           ("color_full_pre", "t_full_pre", True, True)]


def views(name):
    return [(k, str(v)) for k, v in enumerate(SM[f"{name}|views"])]


def frame(name, k, full):
    kw = dict(mode=rt.SHADE_LAMBERT, shadows=True, reflections=True, plane_y=float(SM[f"{name}|plane_y"])) if full else {}
    return rt.make_frame(SM[f"{name}|{k}|rayPos"], SM[f"{name}|{k}|viewInv"], PINV, W, H, **kw)


def buffers(name, k, pre):
    return (SM[f"{name}|{k}|c0"], SM[f"{name}|{k}|t0"]) if pre else (None, None)


def check_all_views(name, render, who, only=None):
    """render(frame, color, t) -> (color, t).  Bit-exact on every view (or those whose label is in `only`)."""
    n = 0
    for k, label in views(name):
        if only is not None and label not in only:
            continue
        for ck, tk, pre, full in RESULTS:
            if only is not None and (pre or full):
                continue  # pre-filled depth, and secondary rays (they start on the surface, inside the boxes), make
                          # the answer tree-dependent as well
            c0, t0 = buffers(name, k, pre)
            col, t = render(frame(name, k, full), c0, t0)
            where = f"{who}, {name} view {k} ({label}), {ck}"
            assert util.same_bits(t, SM[f"{name}|{k}|{tk}"]), f"{where}: depth differs from the reference"
            assert np.array_equal(col, SM[f"{name}|{k}|{ck}"]), f"{where}: colour differs from the reference"
            n += 1
    assert n > 0


def port_render(scene, name):
    full = scene.with_plane(float(SM[f"{name}|plane_y"]))

    def go(f, c0, t0):
        s = full if f.has_plane else scene
        c, t, _, _ = s.render(f.ray_pos, f.view_inv, f.proj_inv, W, H, mode=f.shading_mode, shadows=bool(f.enable_shadows),
                              reflections=bool(f.enable_reflections), color=c0, t=t0)
        return c, t
    return go


# ------------------------------------------------------------------------------------------------------ CPU
@pytest.mark.parametrize("name", ROOT_LEAF)
def test_oracle_every_triangle_loop_is_the_reference_up_to_8_triangles(name):
    s = P.Scene.mesh_brute(SM[f"{name}|vpos4"], SM[f"{name}|indices"])
    check_all_views(name, port_render(s, name), "oracle mesh_brute")


@pytest.mark.parametrize("name", MESHES)
def test_oracle_with_the_reference_tree(name):
    s = P.Scene.mesh(SM[f"{name}|vpos4"], SM[f"{name}|indices_bvh"], SM[f"{name}|nodes"], SM[f"{name}|prim_of_permuted"])
    check_all_views(name, port_render(s, name), "oracle bvh_node")


@pytest.mark.parametrize("name", TREE)
def test_oracle_every_triangle_loop_on_tree_independent_views(name):
    s = P.Scene.mesh_brute(SM[f"{name}|vpos4"], SM[f"{name}|indices"])
    check_all_views(name, port_render(s, name), "oracle mesh_brute", only=("outside",))


@pytest.mark.parametrize("name", ROOT_LEAF)
def test_kernel_source_auto_path_up_to_8_triangles(name):
    s = E.EmulScene.mesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"])  # RT_MESH_AUTO
    check_all_views(name, lambda f, c0, t0: s.render(f, color=c0, t=t0)[:2], "kernel source (host)")


@pytest.mark.parametrize("name", MESHES)
def test_kernel_source_with_the_reference_tree(name):
    s = E.EmulScene.mesh_refbvh(SM[f"{name}|vpos4"], SM[f"{name}|indices_bvh"], SM[f"{name}|nodes"],
                                SM[f"{name}|prim_of_permuted"])
    check_all_views(name, lambda f, c0, t0: s.render(f, color=c0, t=t0)[:2], "kernel source (host), reference tree")


@pytest.mark.parametrize("name", TREE)
def test_kernel_source_own_tree_on_tree_independent_views(name):
    s = E.EmulScene.mesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"])  # RT_MESH_AUTO: our BVH8
    check_all_views(name, lambda f, c0, t0: s.render(f, color=c0, t=t0)[:2], "kernel source (host), own tree",
                    only=("outside",))


def tree_dependence_table(make_scene):
    """Mismatching pixels per (mesh, view label, result) between a path and the reference; not asserted."""
    rows = []
    for name in TREE:
        s = make_scene(name)
        for k, label in views(name):
            for ck, tk, pre, full in RESULTS:
                c0, t0 = buffers(name, k, pre)
                col, t = s(frame(name, k, full), c0, t0)
                bad = int((util.bits(t) != util.bits(SM[f"{name}|{k}|{tk}"])).sum())
                rows.append((name, k, label, ck, bad, W * H))
    return rows


def test_tree_dependent_views_are_counted_not_hidden(tmp_path):
    """Our own tree and the explicit RT_MESH_BRUTE rule on 9+ triangles: where they differ from the reference's tree-
    dependent answers.  Written next to the profiles when run by hand (RT_WRITE_PROFILES=1)."""
    def own(name):
        s = E.EmulScene.mesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"])
        return lambda f, c0, t0: s.render(f, color=c0, t=t0)[:2]

    def boxed(name):
        s = E.EmulScene.mesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"], brute=True)
        return lambda f, c0, t0: s.render(f, color=c0, t=t0)[:2]
    lines = ["# depth pixels that differ from the reference's own frame (48x32), per mesh / view / buffer state",
             "# own = RT_MESH_AUTO (our BVH8), boxed = RT_MESH_BRUTE (mesh-box cull + t <= tFar)",
             f"{'mesh':8s} {'view':>4s} {'label':12s} {'result':15s} {'own':>6s} {'boxed':>6s}"]
    a, b = tree_dependence_table(own), tree_dependence_table(boxed)
    for (name, k, label, ck, bad_a, _), (_, _, _, _, bad_b, _) in zip(a, b):
        lines.append(f"{name:8s} {k:4d} {label:12s} {ck:15s} {bad_a:6d} {bad_b:6d}")
        if label == "outside" and ck == "color":
            assert bad_a == 0 and bad_b == 0, (name, k, ck)
    text = "\n".join(lines) + "\n"
    (tmp_path / "t.txt").write_text(text)
    if os.environ.get("RT_WRITE_PROFILES"):
        with open(os.path.join(os.path.dirname(util.HERE), "profiles", "r2_small_mesh_tree_dependence.txt"), "w") as f:
            f.write(text)


# ------------------------------------------------------------------------------------------------------ GPU
def cuda_render(scene):
    def go(f, c0, t0):
        r = scene.render(f, color=c0, t=t0)
        return r["color"], r["t"]
    return go


@pytest.mark.gpu
@pytest.mark.parametrize("name", ROOT_LEAF)
def test_gpu_auto_path_up_to_8_triangles(name):
    rt.set_math_mode(rt.MATH_EXACT)
    s = rt.TriangleMesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"])
    check_all_views(name, cuda_render(s), "CUDA RT_MESH_AUTO")
    s = rt.TriangleMesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"], accel=rt.MESH_BRUTE)
    check_all_views(name, cuda_render(s), "CUDA RT_MESH_BRUTE")


@pytest.mark.gpu
@pytest.mark.parametrize("name", MESHES)
def test_gpu_with_the_reference_tree(name):
    rt.set_math_mode(rt.MATH_EXACT)
    s = rt.TriangleMeshFromReferenceBVH(SM[f"{name}|vpos4"], SM[f"{name}|indices_bvh"], SM[f"{name}|nodes"],
                                        SM[f"{name}|prim_of_permuted"])
    check_all_views(name, cuda_render(s), "CUDA, reference tree")


@pytest.mark.gpu
@pytest.mark.parametrize("name", TREE)
def test_gpu_own_tree_on_tree_independent_views(name):
    rt.set_math_mode(rt.MATH_EXACT)
    for accel in (rt.MESH_AUTO, rt.MESH_BRUTE, rt.MESH_BVH8):
        s = rt.TriangleMesh(SM[f"{name}|vpos4"], SM[f"{name}|indices"], accel=accel)
        check_all_views(name, cuda_render(s), f"CUDA accel {accel}", only=("outside",))
