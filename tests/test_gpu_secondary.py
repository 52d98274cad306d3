"""GPU: SURVEY 8(f)-1 — the viewer's default look: Lambert + shadow rays + one reflection bounce +
SceneUnion(model, Plane) (raytracing.cpp:33-62, raytracing.hpp:83-97,119-186), and the C++ host mirror driving the
same call sequence as the viewer's frame loop (main.cpp:196-208)."""
import os
import subprocess

import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from tests import util

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = util.Golden()
FULL = [c for c in G.cases if G.case(c)["color_full"] is not None]


@pytest.fixture(autouse=True)
def _defaults():
    rt.set_math_mode(rt.MATH_EXACT)
    rt.set_sched_mode(rt.SCHED_AUTO)
    yield


def _full_frame(c):
    return rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], mode=rt.SHADE_LAMBERT, shadows=True,
                         reflections=True, plane_y=float(c["plane_y"]))


@pytest.mark.parametrize("sched", (rt.SCHED_TILES, rt.SCHED_PERSISTENT))
@pytest.mark.parametrize("key", [c for c in FULL if c.split("|")[0] not in util.MESHES])
def test_sdf_scenes_default_look_bit_exact(key, sched):
    rt.set_sched_mode(sched)
    c = G.case(key)
    s = util.cuda_scene(G, c["scene"])
    r = s.render(_full_frame(c), want_prim=True)
    assert util.same_bits(r["t"], c["t_full"]), key
    assert np.array_equal(r["color"], c["color_full"]), key
    assert (r["prim"] == rt.PRIM_PLANE).any()  # the ground plane is in the frame
    r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], mode=rt.SHADE_LAMBERT,
                               shadows=True))
    assert np.array_equal(r["color"], c["color_shadow"]), key


def test_mesh_default_look_with_the_reference_tree_is_bit_exact():
    """Secondary rays start on the surface and triangle hits are not range-checked, so which triangles
    behind the origin are tested depends on the tree: with the reference's own BVH the frame is reproduced exactly."""
    a = G.scene_arrays("cube_refbvh")
    s = rt.TriangleMeshFromReferenceBVH(a["vpos4"], a["indices_bvh"], a["nodes"], a["prim_of_permuted"])
    for key in [c for c in FULL if c.startswith("cube|")]:
        c = G.case(key)
        r = s.render(_full_frame(c))
        assert util.same_bits(r["t"], c["t_full"]) and np.array_equal(r["color"], c["color_full"])
        r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], mode=rt.SHADE_LAMBERT,
                                   shadows=True))
        assert np.array_equal(r["color"], c["color_shadow"])


# colour pixels of the default look that differ from the reference's frame with OUR tree, counted by running the same
# kernel source on the CPU (tests/host_emul): secondary rays start on the surface, triangle hits are not range-checked
# (SURVEY T6), so a handful of shadow / reflection outcomes depend on the reference's own tree
TREE_DEPENDENT_PIXELS = {"cube|default|128x128": 0, "stanford-bunny|default|160x90": 3, "stanford-bunny|orbit5|96x64": 1,
                         "as1-oc-214|default|96x64": 1}


@pytest.mark.parametrize("key", [c for c in FULL if c.split("|")[0] in util.MESHES])
def test_mesh_default_look_with_our_tree(key):
    """With our own tree primary visibility (depth) is still bit-exact; shading may differ where a secondary ray's
    outcome is tree-dependent in the reference itself.  Those pixels are counted and must be the ones the CPU run of the
    same source found (cube: none, with the tree and with the explicit brute-force rule)."""
    c = G.case(key)
    a = G.scene_arrays(c["scene"])
    accels = (rt.MESH_AUTO, rt.MESH_BRUTE) if c["scene"] == "cube" else (rt.MESH_AUTO,)
    for accel in accels:
        r = rt.TriangleMesh(a["vpos4"], a["indices"], accel=accel).render(_full_frame(c))
        assert util.same_bits(r["t"], c["t_full"]), key
        d = util.compare_frames(r["color"], r["t"], c["color_full"], c["t_full"])
        assert d["mask_mismatch"] == 0
        differ = int((r["color"] != c["color_full"]).sum())
        print(key, "accel", accel, "tree-dependent secondary pixels:", differ, "of", d["n_hit"])
        assert differ <= TREE_DEPENDENT_PIXELS[key], (key, accel, differ, d)


def _write_grid(path, values):
    with open(path, "wb") as f:
        f.write(np.array(values.shape, np.uint32).tobytes())
        f.write(np.ascontiguousarray(values, np.float32).tobytes())


def _write_octree(path, nodes36):
    with open(path, "wb") as f:
        f.write(np.array([nodes36.shape[0]], np.uint32).tobytes())
        f.write(np.ascontiguousarray(nodes36, np.uint8).tobytes())


def _run_example(args, W, H, out):
    exe = os.path.join(ROOT, "triangles_sdf_cpu_raytracing_b200", "hostapi", "example_frame_loop")
    if not os.path.exists(exe):
        pytest.skip("example_frame_loop has not been built (make)")
    p = subprocess.run([exe] + args + [str(out)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    assert "Render Time" in p.stdout
    raw = np.fromfile(out, dtype=np.uint8)
    color = raw[:4 * W * H].view(np.uint32).reshape(H, W)
    t = raw[4 * W * H:].view(np.float32).reshape(H, W)
    return color, t


def test_cpp_frame_loop_reproduces_the_reference_frames(tmp_path):
    """The reference's own class names and call sequence (load scene -> Camera -> clear -> Renderer::draw), compiled
    against hostapi/, produce the reference's frames: files in, framebuffer out."""
    c = G.case("example_grid|default|160x90")
    _write_grid(tmp_path / "g.grid", G.scene_arrays("example_grid")["values"])
    color, t = _run_example([str(tmp_path / "g.grid"), "160", "90", "normal", "0"], 160, 90, tmp_path / "o.raw")
    assert util.same_bits(t, c["t"]) and np.array_equal(color, c["color_normal"])
    color, t = _run_example([str(tmp_path / "g.grid"), "160", "90", "lambert", "1"], 160, 90, tmp_path / "o2.raw")
    assert util.same_bits(t, c["t_full"]) and np.array_equal(color, c["color_full"])
    c = G.case("sdf_6|default|160x90")
    _write_octree(tmp_path / "o.octree", G.scene_arrays("sdf_6")["nodes36"])
    color, t = _run_example([str(tmp_path / "o.octree"), "160", "90", "lambert", "1"], 160, 90, tmp_path / "o3.raw")
    assert util.same_bits(t, c["t_full"]) and np.array_equal(color, c["color_full"])
    # orbit: 5 x Camera::rotate on the host, then the frame of the golden "orbit5" camera
    c = G.case("example_grid|orbit5|96x64")
    exe_args = [str(tmp_path / "g.grid"), "96", "64", "normal", "0"]
    exe = os.path.join(ROOT, "triangles_sdf_cpu_raytracing_b200", "hostapi", "example_frame_loop")
    p = subprocess.run([exe] + exe_args + [str(tmp_path / "o5.raw"), "5"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    raw = np.fromfile(tmp_path / "o5.raw", dtype=np.uint8)
    color = raw[:4 * 96 * 64].view(np.uint32).reshape(64, 96)
    t = raw[4 * 96 * 64:].view(np.float32).reshape(64, 96)
    assert util.same_bits(t, c["t"]) and np.array_equal(color, c["color_normal"])
    # the same loop on a page-locked framebuffer (FrameBuffer::pin + frameIsCleared): the kernel writes host memory
    p = subprocess.run([exe] + exe_args + [str(tmp_path / "o5p.raw"), "5", "pinned"], capture_output=True, text=True,
                       timeout=120)
    assert p.returncode == 0, p.stderr
    assert open(tmp_path / "o5p.raw", "rb").read() == open(tmp_path / "o5.raw", "rb").read()


def test_cpp_frame_loop_obj_mesh(tmp_path):
    """OBJ in -> LoadMeshFromObj + loadAndScale + BVHBuilder::perform on the host -> frame of the golden cube."""
    a = G.scene_arrays("cube")
    # the golden cube is already normalised; write it as an OBJ (normalising it again is the identity up to rounding,
    # so compare against a render of the arrays the C++ loader itself produces)
    with open(tmp_path / "c.obj", "w") as f:
        for v in a["vpos4"]:
            f.write(f"v {v[0]:.9g} {v[1]:.9g} {v[2]:.9g}\n")
        for tri in a["indices"].reshape(-1, 3):
            f.write(f"f {tri[0] + 1} {tri[1] + 1} {tri[2] + 1}\n")
    v, i = rt.host_load_obj(str(tmp_path / "c.obj"), scale=True)
    c = G.case("cube|default|128x128")
    want = rt.TriangleMesh(v, i).render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], 128, 128))
    color, t = _run_example([str(tmp_path / "c.obj"), "128", "128", "normal", "0"], 128, 128, tmp_path / "o.raw")
    assert util.same_bits(t, want["t"]) and np.array_equal(color, want["color"])
    assert np.isfinite(t).mean() > 0.3


def test_cpp_frame_loop_ppm_sink(tmp_path):
    """SURVEY 8(f)-4, the step after the path: frames leave as binary PPM files (one per orbit frame) instead of an
    SDL blit; the RGB bytes are the colour buffer's."""
    exe = os.path.join(ROOT, "triangles_sdf_cpu_raytracing_b200", "hostapi", "example_frame_loop")
    if not os.path.exists(exe):
        pytest.skip("example_frame_loop has not been built (make)")
    _write_grid(tmp_path / "g.grid", G.scene_arrays("example_grid")["values"])
    p = subprocess.run([exe, str(tmp_path / "g.grid"), "96", "64", "normal", "0", str(tmp_path / "f.ppm"), "5"],
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    names = sorted(n for n in os.listdir(tmp_path) if n.endswith(".ppm"))
    assert names == [f"f_{k:03d}.ppm" for k in range(6)]
    raw = open(tmp_path / "f_005.ppm", "rb").read()
    head = b"P6\n96 64\n255\n"
    assert raw.startswith(head) and len(raw) == len(head) + 96 * 64 * 3
    rgb = np.frombuffer(raw[len(head):], np.uint8).reshape(64, 96, 3)
    want = G.case("example_grid|orbit5|96x64")["color_normal"].view(np.uint8).reshape(64, 96, 4)[..., :3]
    assert np.array_equal(rgb, want)
