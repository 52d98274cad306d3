"""GPU: the CUDA path, called through the C ABI, against (a) the golden frames the reference itself rendered and
(b) the plain-C oracle on the same seeded inputs.  RT_MATH_EXACT must be bit-identical; RT_MATH_FAST must meet the
north-star tolerances (hit mask exact bar counted pixels, depth 1e-4 relative, RGB within 1/255)."""
import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from oracle import port_oracle as P
from tests import util

pytestmark = pytest.mark.gpu

G = util.Golden()
SDF_CASES = [c for c in G.cases if c.split("|")[0] in ("example_grid", "sdf_5", "sdf_6")]
MESH_CASES = [c for c in G.cases if c.split("|")[0] in util.MESHES]
_scene_cache = {}


def assert_prim_mismatches_are_exact_ties(c, prim, want_prim, t, limit=0.02):
    """Prim ids may differ from the reference's only where two triangles give the bit-identical closest t (shared
    edges, duplicated geometry): the winner then depends on the visit order, i.e. on the tree.
    Every such pixel is re-checked with the oracle's single-triangle test; they are counted and bounded."""
    a = G.scene_arrays(c["scene"])
    v = a["vpos4"].reshape(-1, 4)
    idx = a["indices"].reshape(-1, 3)
    ys, xs = np.nonzero(prim != want_prim)
    assert ys.size <= limit * prim.size, f"{ys.size} prim-id mismatches"
    for row, x in zip(ys, xs):
        assert prim[row, x] != rt.PRIM_MISS and want_prim[row, x] != rt.PRIM_MISS
        d = P.eye_ray(c["rayPos"], c["viewInv"], c["projInv"], prim.shape[1], prim.shape[0], int(x),
                      int(prim.shape[0] - 1 - row))
        tri = v[idx[prim[row, x]], :3] / v[idx[prim[row, x]], 3:4]
        hit, tt, _ = P.triangle(c["rayPos"], d, tri.astype(np.float32))
        assert hit and util.same_bits(tt, t[row, x]), f"pixel ({x},{row}): prim {prim[row, x]} is not a tie"
    return ys.size


def scene(name, **kw):
    key = (name, tuple(sorted(kw.items())))
    if key not in _scene_cache:
        _scene_cache[key] = util.cuda_scene(G, name, **kw)
    return _scene_cache[key]


@pytest.fixture(autouse=True)
def _defaults():
    rt.set_math_mode(rt.MATH_EXACT)
    rt.set_sched_mode(rt.SCHED_AUTO)
    yield
    rt.set_math_mode(rt.MATH_EXACT)
    rt.set_sched_mode(rt.SCHED_AUTO)


def _frame(c, **kw):
    return rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], **kw)


@pytest.mark.parametrize("sched", (rt.SCHED_TILES, rt.SCHED_PERSISTENT, rt.SCHED_COMPACT))
@pytest.mark.parametrize("key", SDF_CASES)
def test_sdf_scenes_bit_exact_vs_reference(key, sched):
    rt.set_sched_mode(sched)
    c = G.case(key)
    s = scene(c["scene"])
    for mode, name in ((rt.SHADE_NORMAL, "color_normal"), (rt.SHADE_LAMBERT, "color_lambert"),
                       (rt.SHADE_COLOR, "color_color")):
        r = s.render(_frame(c, mode=mode))
        assert util.same_bits(r["t"], c["t"]), f"{key}: depth differs on {(util.bits(r['t']) != util.bits(c['t'])).sum()} px"
        assert np.array_equal(r["color"], c[name]), f"{key}: {name} differs"
    # in/out depth buffer
    t0 = np.full((c["H"], c["W"]), 2.2, np.float32)
    c0 = np.full((c["H"], c["W"]), 0x11223344, np.uint32)
    r = s.render(_frame(c), color=c0, t=t0)
    assert util.same_bits(r["t"], c["t_prefilled"]) and np.array_equal(r["color"], c["color_prefilled"])


def test_grid_brick_layout_is_bit_exact():
    for key in [c for c in SDF_CASES if c.startswith("example_grid")]:
        c = G.case(key)
        r = scene("example_grid", layout=rt.GRID_BRICKS).render(_frame(c))
        assert util.same_bits(r["t"], c["t"]) and np.array_equal(r["color"], c["color_normal"])


@pytest.mark.parametrize("accel", (rt.MESH_BVH8, rt.MESH_BRUTE, rt.MESH_LBVH_GPU))
@pytest.mark.parametrize("key", MESH_CASES)
def test_mesh_scenes_vs_reference(key, accel):
    c = G.case(key)
    if accel == rt.MESH_BRUTE and c["scene"] != "cube":
        pytest.skip("brute force is for small meshes")
    if accel == rt.MESH_LBVH_GPU and c["scene"] == "cube":
        pytest.skip("12 triangles: nothing for the device builder to build")
    s = scene(c["scene"], accel=accel)
    r = s.render(_frame(c), want_prim=True)
    # closest hit does not depend on the tree: depth and colour are bit-identical, prim ids may differ on exact ties
    assert util.same_bits(r["t"], c["t"]), f"{key}: depth differs on {(util.bits(r['t']) != util.bits(c['t'])).sum()} px"
    assert_prim_mismatches_are_exact_ties(c, r["prim"], c["prim"], c["t"])
    same = r["prim"] == c["prim"]  # on a tie between two triangles the normal (colour) is the other triangle's
    assert np.array_equal(r["color"][same], c["color_normal"][same])
    assert util.rgb_max_diff(r["color"][~same] >> 24, c["color_normal"][~same] >> 24) == 0  # alpha: hit everywhere
    r = s.render(_frame(c, mode=rt.SHADE_LAMBERT))
    assert np.array_equal(r["color"][same], c["color_lambert"][same])


def test_device_built_bvh_structure_and_large_mesh():
    """RT_MESH_LBVH_GPU: the tree built on the device renders the same closest hits as the host-built
    one on the 1.1 M-triangle stand-in, and reports its size / build time."""
    a = G.scene_arrays("stanford-bunny_s2")
    host = rt.TriangleMesh(a["vpos4"], a["indices"], accel=rt.MESH_BVH8)
    dev = rt.TriangleMesh(a["vpos4"], a["indices"], accel=rt.MESH_LBVH_GPU)
    st = dev.mesh_stats()
    assert st["n_tris"] == a["indices"].size // 3 and 0 < st["n_nodes"] < st["n_tris"] // 2
    assert 0.0 < st["device_build_ms"] < 200.0
    assert host.mesh_stats()["device_build_ms"] == 0.0
    assert np.allclose(dev.info()["bbox_min"], host.info()["bbox_min"]) and np.allclose(dev.info()["bbox_max"], host.info()["bbox_max"])
    c = G.case("stanford-bunny|default|160x90")
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], 1920, 1080)
    rh, rd = host.render(f, want_prim=True), dev.render(f, want_prim=True)
    assert util.same_bits(rh["t"], rd["t"])
    same = rh["prim"] == rd["prim"]
    assert (~same).sum() <= 0.001 * same.size          # exact-t ties between neighbouring triangles only
    assert np.array_equal(rh["color"][same], rd["color"][same])
    # malformed input is rejected, not read out of bounds
    bad = a["indices"].copy()
    bad[5] = a["vpos4"].shape[0] + 7
    with pytest.raises(rt.RtError):
        rt.TriangleMesh(a["vpos4"], bad, accel=rt.MESH_LBVH_GPU)
    host.destroy()
    dev.destroy()


def test_reference_built_bvh_is_traversed_identically():
    a = G.scene_arrays("cube_refbvh")
    s = rt.TriangleMeshFromReferenceBVH(a["vpos4"], a["indices_bvh"], a["nodes"], a["prim_of_permuted"])
    for key in [c for c in MESH_CASES if c.startswith("cube|")]:
        c = G.case(key)
        r = s.render(_frame(c), want_prim=True)
        assert util.same_bits(r["t"], c["t"]) and np.array_equal(r["color"], c["color_normal"])
        assert np.array_equal(r["prim"], c["prim"])
        # with the reference's own tree even the tree-dependent tFar quirk is reproduced
        t0 = np.full((c["H"], c["W"]), 2.2, np.float32)
        c0 = np.full((c["H"], c["W"]), 0x11223344, np.uint32)
        r = s.render(_frame(c), color=c0, t=t0)
        assert util.same_bits(r["t"], c["t_prefilled"]) and np.array_equal(r["color"], c["color_prefilled"])


@pytest.mark.parametrize("name", ("example_grid", "sdf_6", "stanford-bunny", "cube"))
def test_single_rays_match_reference(name):
    """IScene::intersect on seeded rays incl. exactly axis-parallel ones (inf / NaN slab arithmetic)."""
    k = G.kats
    hit, t, n, _ = scene(name).intersect(k[f"ray_{name}_o"], k[f"ray_{name}_d"])
    want = k[f"ray_{name}_hit"].astype(bool)
    assert np.array_equal(hit, want)
    assert util.same_bits(t[want], k[f"ray_{name}_t"][want])
    if name == "cube":  # a ray through a face's diagonal ties between its two triangles; their normals are the same
        assert np.array_equal(n[want], k[f"ray_{name}_n"][want])  # vector up to the sign of a zero (which one wins is the tree's)
    else:
        assert util.same_bits(n[want], k[f"ray_{name}_n"][want])


@pytest.mark.parametrize("name,W,H", (("example_grid", 480, 270), ("sdf_6", 480, 270), ("stanford-bunny", 320, 180),
                                      ("cube", 256, 256)))
def test_larger_frames_vs_c_oracle(name, W, H):
    """Same seeded inputs through the plain-C oracle at a size it finishes in seconds; counters must agree too."""
    c = G.case([k for k in G.cases if k.startswith(name + "|default")][0])
    # the projection only depends on the aspect ratio: reuse the golden camera when it matches
    assert abs(W / H - c["W"] / c["H"]) < 1e-6
    ps = util.port_scene(G, name)
    want_c, want_t, want_p, want_cnt = ps.render(c["rayPos"], c["viewInv"], c["projInv"], W, H, want_counters=True)
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    r = scene(name).render(f, want_prim=True, want_counters=True)
    assert util.same_bits(r["t"], want_t) and np.array_equal(r["color"], want_c)
    if name in ("example_grid", "sdf_6"):
        assert np.array_equal(r["prim"], want_p)
        cn = r["counters"]
        assert cn["rays"] == want_cnt["rays"] and cn["hits"] == want_cnt["hits"]
        assert cn["grid_samples"] == want_cnt["gridSdf"]
        assert cn["oct_inner"] == want_cnt["octInner"] and cn["oct_leaf"] == want_cnt["octLeaf"]
        assert cn["oct_leaf_samples"] == want_cnt["octLeafSdf"]
    else:
        cc = dict(c, W=W, H=H)
        assert_prim_mismatches_are_exact_ties(cc, r["prim"], want_p, want_t)


@pytest.mark.parametrize("key", [c for c in G.cases if "|default|" in c or "|orbit5|" in c])
def test_fast_math_meets_north_star_tolerances(key):
    rt.set_math_mode(rt.MATH_FAST)
    c = G.case(key)
    r = scene(c["scene"]).render(_frame(c))
    d = util.compare_frames(r["color"], r["t"], c["color_normal"], c["t"])
    n = c["W"] * c["H"]
    # grazing / edge pixels are counted and bounded, not hidden
    assert d["mask_mismatch"] <= 0.002 * n, d
    assert d["depth_outliers"] <= 0.005 * n, d
    assert d["rgb_outliers"] <= 0.005 * n, d


def test_screen_space_bands_compose_the_full_frame():
    """Multi-GPU partition (interleaved row bands): the union of all bands equals the single-call frame."""
    c = G.case("sdf_6|default|160x90")
    s = scene("sdf_6")
    full = s.render(_frame(c))
    for band_rows, count in ((8, 2), (16, 4), (7, 3), (32, 8)):
        color = np.zeros((c["H"], c["W"]), np.uint32)
        t = np.full((c["H"], c["W"]), np.inf, np.float32)
        for i in range(count):
            r = s.render(_frame(c, bands=(band_rows, count, i)), color=color, t=t)
            color, t = r["color"], r["t"]
        assert util.same_bits(t, full["t"]) and np.array_equal(color, full["color"])


def test_edge_cases():
    # 1x1 and ragged sizes (not multiples of the 8x4 warp tile)
    c = G.case("example_grid|default|160x90")
    s = scene("example_grid")
    ps = util.port_scene(G, "example_grid")
    for W, H in ((1, 1), (3, 5), (33, 9), (8, 4)):
        pinv = c["projInv"]  # aspect differs: still a valid projection, both sides get the same matrix
        wc, wt, _, _ = ps.render(c["rayPos"], c["viewInv"], pinv, W, H)
        r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], pinv, W, H))
        assert util.same_bits(r["t"], wt) and np.array_equal(r["color"], wc)
    # degenerate scenes
    one = np.zeros((1, 36), np.uint8)  # a single all-zero leaf: isEmpty
    r = rt.SDFOctree(one).render(_frame(c))
    assert not np.isfinite(r["t"]).any()
    with pytest.raises(rt.RtError):
        bad = np.zeros((2, 36), np.uint8)
        bad[0, 32:36] = np.array([5], np.uint32).view(np.uint8)  # children 5..12 out of range
        rt.SDFOctree(bad)
    with pytest.raises(rt.RtError):
        rt.TriangleMesh(np.zeros((3, 4), np.float32), np.array([0, 1, 7], np.uint32))
    with pytest.raises(rt.RtError):
        rt.SDFGrid(np.zeros((1, 4, 4), np.float32))
    # a degenerate triangle never hits (|det| < 1e-8) and does not poison its neighbours
    v = np.array([[0, 0, 0, 1], [1, 0, 0, 1], [2, 0, 0, 1], [-1, -1, 0, 1], [1, -1, 0, 1], [0, 1, 0, 1]], np.float32)
    m = rt.TriangleMesh(v, np.array([0, 1, 2, 3, 4, 5], np.uint32))
    hit, t, n, prim = m.intersect(np.array([[0, 0, 2.5]], np.float32), np.array([[0, 0, -1]], np.float32))
    assert hit[0] and prim[0] == 1 and t[0] == np.float32(2.5)


@pytest.mark.parametrize("name", ("sdf_6", "sdf_5", "example_grid", "cube"))
def test_random_rays_vs_c_oracle(name):
    """IScene::intersect on 40k seeded rays per scene: from outside, from INSIDE the model's box (every box entry
    clamps to tNear: sort8 ties), axis-parallel and through exact cell/octant corners."""
    rng = np.random.default_rng(7)
    n = 40000
    o = rng.uniform(-1.6, 1.6, (n, 3)).astype(np.float32)
    tgt = rng.uniform(-0.7, 0.7, (n, 3)).astype(np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    # rays through exact dyadic points along exact diagonals / axes: ties between sibling boxes
    k = 4000
    o[:k] = np.float32(2.0) * np.sign(rng.uniform(-1, 1, (k, 3))).astype(np.float32)
    corner = (rng.integers(-8, 9, (k, 3)) / 8.0).astype(np.float32)
    dd = corner - o[:k]
    d[:k] = (dd / np.linalg.norm(dd, axis=1, keepdims=True)).astype(np.float32)
    o[k:2 * k, 1:] = (rng.integers(-8, 9, (k, 2)) / 8.0).astype(np.float32)
    o[k:2 * k, 0] = 2.5
    d[k:2 * k] = np.array([-1.0, 0.0, 0.0], np.float32)
    ps = util.port_scene(G, name)
    # (cube: rays that start inside a mesh are tree-dependent; the explicit brute-force rule is the oracle loop's)
    hit, t, nrm, prim = scene(name, **({"accel": rt.MESH_BRUTE} if name == "cube" else {})).intersect(o, d)
    bad = 0
    for i in range(n):
        h, tt, nn, pp = ps.intersect(o[i], d[i])
        assert h == bool(hit[i]), f"ray {i}: hit mask"
        if h:
            assert util.same_bits(tt, t[i]), f"ray {i}: t {tt} vs {t[i]}"
            if name != "cube":
                assert pp == int(prim[i]), f"ray {i}: prim"
                assert util.same_bits(nn, nrm[i]) or (np.isnan(nn).any() and np.isnan(nrm[i]).any()), f"ray {i}: normal"
            elif pp != int(prim[i]):
                bad += 1  # exact ties between the two triangles of a face
    assert bad <= 0.02 * n


def test_fused_clear_equals_clear_then_draw():
    """rt_frame.clear: FrameBuffer::clear() fused into the render launch; dirty buffers in, the reference's frame out."""
    for key in ("example_grid|default|160x90", "sdf_6|default|160x90", "cube|default|128x128"):
        c = G.case(key)
        dirty_c = np.full((c["H"], c["W"]), 0xDEADBEEF, np.uint32)
        dirty_t = np.full((c["H"], c["W"]), 0.5, np.float32)
        r = scene(c["scene"]).render(_frame(c, clear=True), color=dirty_c, t=dirty_t, want_prim=True)
        assert util.same_bits(r["t"], c["t"]) and np.array_equal(r["color"], c["color_normal"])
        assert (r["prim"][~np.isfinite(c["t"])] == rt.PRIM_MISS).all()


def test_staged_peer_store_kernel_is_bit_exact():
    """The multi-GPU store path (tile parked in shared memory, 128-byte TMA bulk stores) forced on a local framebuffer:
    RT_STAGE_STORES is read once at load, so the check runs in its own process."""
    import os
    import subprocess
    import sys
    code = """
import numpy as np, torch
import triangles_sdf_cpu_raytracing_b200 as rt
from tests import util
g = util.Golden()
for key in ("example_grid|default|160x90", "sdf_6|default|160x90", "stanford-bunny|default|160x90", "sdf_6|orbit23tilt|96x64"):
    c = g.case(key)
    s = util.cuda_scene(g, c["scene"])
    W, H = c["W"], c["H"]
    color = torch.full((H * W,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    t = torch.full((H * W,), 0.25, dtype=torch.float32, device="cuda")
    for bands in (None, (8, 3, 1)):
        f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True, bands=bands)
        s.render_device(f, color.data_ptr(), None, t.data_ptr())
    torch.cuda.synchronize()
    assert util.same_bits(t.cpu().numpy().reshape(H, W), c["t"]), key
    assert np.array_equal(color.cpu().numpy().view(np.uint32).reshape(H, W), c["color_normal"]), key
print("staged ok")
"""
    env = dict(os.environ, RT_STAGE_STORES="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "staged ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["0", "1", "2"])
def test_host_zero_copy_is_bit_exact(mode):
    """rt_render_host on a pinned host framebuffer, every transfer mode (RT_HOST_ZEROCOPY: 0 = pipelined copies,
    1 = plain stores into host memory, 2 = staged 128-byte bulk stores, the default).  Dirty buffers, fused clear: every
    pixel must come back as the reference's.  The knob is read once at load, so the check runs in its own process."""
    import os
    import subprocess
    import sys
    code = """
import ctypes as C
import numpy as np, torch
import triangles_sdf_cpu_raytracing_b200 as rt
from triangles_sdf_cpu_raytracing_b200 import _lib as L
from tests import util
g = util.Golden()
for key in ("example_grid|default|160x90", "sdf_6|default|160x90", "stanford-bunny|default|160x90", "sdf_6|orbit23tilt|96x64",
            "cube|default|128x128"):
    c = g.case(key)
    s = util.cuda_scene(g, c["scene"])
    W, H = c["W"], c["H"]
    color = torch.full((H * W,), 0x5A5A5A5A, dtype=torch.int32).pin_memory()
    t = torch.full((H * W,), 0.25, dtype=torch.float32).pin_memory()
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True)
    dms, tms = C.c_float(), C.c_float()
    for _ in range(2):
        L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(color.data_ptr()), C.c_void_p(t.data_ptr()), None,
                                       C.byref(dms), C.byref(tms), None))
    assert tms.value > 0.0
    assert util.same_bits(t.numpy().reshape(H, W), c["t"]), key
    assert np.array_equal(color.numpy().view(np.uint32).reshape(H, W), c["color_normal"]), key
    # the reference's in/out contract (raytracing.cpp:90-94): left half cleared on the host, right half already holds a
    # nearer surface that must survive
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    cn, tn = color.numpy().view(np.uint32).reshape(H, W), t.numpy().reshape(H, W)
    cn[:, :W // 2] = 0
    tn[:, :W // 2] = np.inf
    cn[:, W // 2:] = 7
    tn[:, W // 2:] = 1e-3
    L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(color.data_ptr()), C.c_void_p(t.data_ptr()), None,
                                   C.byref(dms), C.byref(tms), None))
    assert util.same_bits(tn[:, :W // 2], c["t"][:, :W // 2]), key
    assert np.array_equal(cn[:, :W // 2], c["color_normal"][:, :W // 2]), key
    # (also for the 12-triangle cube's brute-force kernel: its mesh-bounding-box cull stands in for the tree's boxes)
    assert np.all(cn[:, W // 2:] == 7) and np.all(tn[:, W // 2:] == np.float32(1e-3)), key
print("zero-copy ok")
"""
    env = dict(os.environ, RT_HOST_ZEROCOPY=mode)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "zero-copy ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_pinned_framebuffer_and_registered_arrays():
    """The page-locked framebuffer API (rt_pinned_alloc / rt_pinned_register, include/rt_b200.h): PinnedFrameBuffer +
    Scene.render_into with the fused clear (kernel stores into host memory) and with the reference's in/out contract
    (copies), and numpy arrays registered in place."""
    import ctypes as C
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    for key in ("example_grid|default|160x90", "sdf_6|default|160x90", "stanford-bunny|default|160x90"):
        c = G.case(key)
        s = util.cuda_scene(G, c["scene"])
        W, H = c["W"], c["H"]
        fb = rt.PinnedFrameBuffer(W, H)
        fb.color[...] = 0x5A5A5A5A
        fb.t[...] = 0.25
        dms, tms = s.render_into(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True), fb)
        assert tms > 0.0
        assert util.same_bits(fb.t, c["t"]) and np.array_equal(fb.color, c["color_normal"]), key
        fb.clear()
        s.render_into(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), fb)
        assert util.same_bits(fb.t, c["t"]) and np.array_equal(fb.color, c["color_normal"]), key
        # in/out: a nearer depth already in the buffer wins (raytracing.cpp:90-94)
        fb.color[...] = 7
        fb.t[...] = 1e-3
        s.render_into(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), fb)
        assert np.all(fb.color == 7) and np.all(fb.t == np.float32(1e-3)), key
        fb.free()
        def own_pages(dtype):  # two registered ranges must not share a page: page-aligned, padded to whole pages
            raw = np.empty(H * W * 4 + 8192, np.uint8)
            off = (-raw.ctypes.data) % 4096
            return raw[off:off + H * W * 4].view(dtype).reshape(H, W)
        color, t = own_pages(np.uint32), own_pages(np.float32)
        color[...] = 0x5A5A5A5A
        t[...] = 0.25
        L.check(L.lib().rt_pinned_register(C.c_void_p(color.ctypes.data), color.nbytes))
        L.check(L.lib().rt_pinned_register(C.c_void_p(t.ctypes.data), t.nbytes))
        try:
            f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True)
            L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(color.ctypes.data), C.c_void_p(t.ctypes.data),
                                           None, None, None, None))
        finally:
            L.check(L.lib().rt_pinned_unregister(C.c_void_p(color.ctypes.data)))
            L.check(L.lib().rt_pinned_unregister(C.c_void_p(t.ctypes.data)))
        assert util.same_bits(t, c["t"]) and np.array_equal(color, c["color_normal"]), key


def test_brute_force_path_box_cull_and_depth_bound():
    """The explicit RT_MESH_BRUTE rule on the cube (BRUTE_BOXED, rt_scene.h): looking away from the cube draws nothing, a
    nearer depth already in the framebuffer survives, eyes inside the cube see what the reference sees (hits behind the
    origin included).  Against the oracle's every-triangle loop, which tests/test_host_emul.py pins to the compiled
    reference on the same situations for THIS mesh; other small meshes are tests/test_small_meshes.py's."""
    a = G.scene_arrays("cube")
    s = rt.TriangleMesh(a["vpos4"], a["indices"], accel=rt.MESH_BRUTE)
    ps = P.Scene.mesh_brute(a["vpos4"], a["indices"])
    W, H = 64, 40
    pinv = rt.host_projection(W, H)[1]
    rng = np.random.default_rng(3)
    drew_nothing = 0
    for radii in ((1.8, 2.5, 4.0), (1.05, 1.2), (0.2, 0.5, 0.9)):
        for prefill in (False, True):
            for away in (False, True):
                for k in range(4):
                    v = rng.normal(size=3)
                    v /= np.linalg.norm(v)
                    r = float(rng.choice(radii))
                    target = 2.5 * r * v + rng.uniform(-0.5, 0.5, 3) if away else rng.uniform(-0.3, 0.3, size=3)
                    pos, _view, vinv = rt.host_camera(eye=tuple(float(x) for x in r * v),
                                                      target=tuple(float(x) for x in target), lock_up=bool(k % 2))
                    t0 = (rng.choice(np.array([np.inf, 0.5, 1.2, 2.0, 2.6, 3.5], np.float32), size=(H, W)).astype(np.float32)
                          if prefill else np.full((H, W), np.inf, np.float32))
                    c0 = rng.integers(0, 2**32, size=(H, W), dtype=np.uint32)
                    got = s.render(rt.make_frame(pos, vinv, pinv, W, H), color=c0, t=t0)
                    pc, pt, _, _ = ps.render(pos, vinv, pinv, W, H, color=c0, t=t0)
                    where = f"radii {radii} prefill {prefill} away {away} camera {k}"
                    assert util.same_bits(got["t"], pt), where
                    assert np.array_equal(got["color"], pc), where  # same triangle order, same strict '>' on ties
                    if away and radii[0] > 1.7 and not prefill:
                        drew_nothing += int(np.isinf(got["t"]).all())
    assert drew_nothing == 4
    s.destroy()


def _octree_with_shared_leaves():
    """Root (children 1..8); children 1 and 6 are inner nodes that SHARE one block of 8 leaves (9..16): a file with
    de-duplicated leaves, which the reference's loader and intersectNode accept (octree_raytracing.cpp:8-16,166-202)."""
    n = 17
    rec = np.zeros((n, 9), np.uint32)
    val = rec[:, :8].view(np.float32)
    rec[0, 8] = 1
    val[1:9] = 1000.0            # empty-outside leaves ...
    rec[1, 8] = 9                # ... except two inner nodes pointing at the same block
    rec[6, 8] = 9
    val[1] = 0.0
    val[6] = 0.0
    rng = np.random.default_rng(11)
    val[9:17] = rng.uniform(-0.2, 0.3, (8, 8)).astype(np.float32)   # surface leaves
    val[12] = 1000.0             # one culled leaf inside the shared block
    return rec.view(np.uint8).reshape(n, 36)


def test_octree_with_shared_leaves_loads_and_matches_the_oracle():
    from oracle import port_oracle as P
    nodes = _octree_with_shared_leaves()
    s = rt.SDFOctree(nodes)
    ps = P.Scene.octree(nodes)
    c = G.case("sdf_6|default|160x90")
    for cam in ("sdf_6|default|160x90", "sdf_6|orbit23tilt|96x64"):
        c = G.case(cam)
        want_c, want_t, _, want_cnt = ps.render(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], want_counters=True)
        r = s.render(_frame(c), want_counters=True)
        assert util.same_bits(r["t"], want_t) and np.array_equal(r["color"], want_c), cam
        assert r["counters"]["oct_inner"] == want_cnt["octInner"] and r["counters"]["oct_leaf"] == want_cnt["octLeaf"]
        r2 = s.render(_frame(c))  # the plain flavour drops culled children when it builds its lists
        assert util.same_bits(r2["t"], want_t) and np.array_equal(r2["color"], want_c), cam
    assert np.isfinite(want_t).any()
    # an inner node referenced twice is how a cycle would look: rejected
    bad = nodes.copy().view(np.uint32).reshape(-1, 9)
    bad[9, 8] = 1  # leaf 9 becomes an inner node pointing back at the root's children
    with pytest.raises(rt.RtError):
        rt.SDFOctree(bad.view(np.uint8).reshape(-1, 36))


def test_pinned_buffers_the_bulk_copies_cannot_address_fall_back_to_the_copy_pipeline():
    """rt_render_host takes the kernel-stores-to-host route only for 16-byte aligned, fully page-locked buffers
    (include/rt_b200.h); an offset or partly registered range silently takes the copies and still returns the frame."""
    import ctypes as C
    import torch
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    c = G.case("example_grid|default|160x90")
    s = scene("example_grid")
    W, H = c["W"], c["H"]
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True)
    big_c = torch.zeros(W * H + 64, dtype=torch.int32).pin_memory()
    big_t = torch.zeros(W * H + 64, dtype=torch.float32).pin_memory()
    for off in (1, 3):  # 4- and 12-byte offsets: not 16-byte aligned
        col, t = big_c[off:off + W * H], big_t[off:off + W * H]
        col.fill_(0x5A5A5A5A)
        t.fill_(0.25)
        L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(col.data_ptr()), C.c_void_p(t.data_ptr()), None,
                                       None, None, None))
        assert util.same_bits(t.numpy().reshape(H, W), c["t"]) and np.array_equal(col.numpy().view(np.uint32).reshape(H, W), c["color_normal"]), off
    # Only the first half of each buffer is registered.  The kernel must not be pointed at it (an illegal address inside a
    # kernel is a sticky fault that kills the context); the copy pipeline either copes or the runtime refuses the copy
    # with an ordinary error (cudaMemcpyAsync rejects a range that is half pinned, half pageable), and the library goes on.
    def pages(dtype):
        raw = np.empty(H * W * 4 + 8192, np.uint8)
        o = (-raw.ctypes.data) % 4096
        return raw[o:o + H * W * 4].view(dtype).reshape(H, W)
    col, t = pages(np.uint32), pages(np.float32)
    half = (H * W * 4 // 2) // 4096 * 4096
    L.check(L.lib().rt_pinned_register(C.c_void_p(col.ctypes.data), half))
    L.check(L.lib().rt_pinned_register(C.c_void_p(t.ctypes.data), half))
    try:
        col[...] = 7
        t[...] = 0.5
        rc = L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(col.ctypes.data), C.c_void_p(t.ctypes.data), None,
                                    None, None, None)
    finally:
        L.check(L.lib().rt_pinned_unregister(C.c_void_p(col.ctypes.data)))
        L.check(L.lib().rt_pinned_unregister(C.c_void_p(t.ctypes.data)))
    if rc == 0:
        assert util.same_bits(t, c["t"]) and np.array_equal(col, c["color_normal"])
    else:
        assert rc == -2 and b"cudaMemcpy" in L.lib().rt_last_error()
    r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H))  # the context is alive and well
    assert util.same_bits(r["t"], c["t"]) and np.array_equal(r["color"], c["color_normal"])


def test_grid_with_non_finite_samples_takes_the_literal_path():
    """The linear-layout sampler relies on finite samples (csrc/rt_device.cuh: grid_sdf); a grid holding infinities is
    stored bricked and sampled literally, and still equals the oracle."""
    from oracle import port_oracle as P
    c = G.case("example_grid|default|160x90")
    v = G.scene_arrays("example_grid")["values"].copy()
    far = v > 0.9
    assert far.any()
    v[far] = np.inf  # "far away" markers
    want_c, want_t, _, _ = P.Scene.grid(v).render(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"])
    r = rt.SDFGrid(v).render(_frame(c))
    assert util.same_bits(r["t"], want_t) and np.array_equal(r["color"], want_c)
    assert np.isfinite(want_t).any()
