"""CPU: the product's per-ray CUDA source (csrc/rt_device.cuh + csrc/rt_shade.cuh) compiled as plain C++
(tests/host_emul/) against the golden frames of the reference and against the plain-C oracle on seeded cameras.

The GPU parity tests check the kernels on a few dozen frames; this file runs the same source text on the CPU over
hundreds of cameras (eyes inside the [-1,1]^3 box and inside meshes included), every shading mode, secondary rays, the
ground plane and the in/out depth buffer, bit for bit.  It covers the arithmetic and control flow of the per-ray code;
what only exists on the device (tiles, bands, shared-memory staging, TMA stores, warp votes) is the GPU tests' job."""
import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from oracle import port_oracle as P
from tests import util
from tests import host_emul as E

G = util.Golden()
# rt_counters name -> the oracle's name
COUNTERS = {"grid": {"rays": "rays", "hits": "hits", "grid_samples": "gridSdf"},
            "octree": {"rays": "rays", "hits": "hits", "oct_inner": "octInner", "oct_leaf": "octLeaf",
                       "oct_leaf_samples": "octLeafSdf"}}


@pytest.fixture(scope="module", autouse=True)
def _built():
    if E.cuda_include() is None:
        pytest.skip("no CUDA headers on this machine (float4 / make_float4 come from cuda_runtime.h)")
    E.build()


def emul_scene(name, **kw):
    a = G.scene_arrays(name)
    if "values" in a:
        return E.EmulScene.grid(a["values"]), "grid"
    if "nodes36" in a:
        return E.EmulScene.octree(a["nodes36"]), "octree"
    return E.EmulScene.mesh(a["vpos4"], a["indices"], **kw), "mesh"


def seeded_cameras(n, seed, radii=(0.2, 0.35, 0.9, 1.0, 1.6, 2.5, 4.0)):
    """Eyes on spheres of the given radii (< 1.74: inside the bounding sphere of the [-1,1]^3 box), off-centre targets,
    locked / free up vector, mouse rotations: (rayPos, viewInv) through the host mirror of Camera (camera.cpp)."""
    rng = np.random.default_rng(seed)
    for k in range(n):
        v = rng.normal(size=3)
        eye = float(rng.choice(radii)) * v / np.linalg.norm(v)
        target = rng.uniform(-0.3, 0.3, size=3)
        pos, _view, vinv = rt.host_camera(eye=tuple(float(x) for x in eye), target=tuple(float(x) for x in target),
                                          lock_up=bool(k % 2), n_rotate=int(rng.integers(0, 4)),
                                          dx=float(rng.uniform(-80, 80)), dy=float(rng.uniform(-40, 40)))
        yield k, pos, vinv


@pytest.mark.parametrize("key", G.cases)
def test_device_source_on_cpu_reproduces_the_golden_frames(key):
    c = G.case(key)
    s, kind = emul_scene(c["scene"])
    W, H = c["W"], c["H"]
    for mode, name in ((rt.SHADE_NORMAL, "color_normal"), (rt.SHADE_LAMBERT, "color_lambert"), (rt.SHADE_COLOR, "color_color")):
        col, t, prim, _ = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=mode))
        assert util.same_bits(t, c["t"]), f"{key}: depth differs"
        # the plain flavour (no workload counters: what the timed launches run; the octree filters its child lists there)
        col2, t2, prim2, _ = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=mode), count=False)
        assert util.same_bits(t2, t) and np.array_equal(col2, col) and np.array_equal(prim2, prim), f"{key}: flavours differ"
        if kind == "mesh":  # exact-t ties between two triangles: the winner depends on the tree
            assert (col != c[name]).sum() <= max(2, 1e-3 * W * H), f"{key}: {name} differs beyond tie pixels"
        else:
            assert np.array_equal(col, c[name]), f"{key}: {name} differs"
    if "color_full" in c and c["color_full"] is not None:
        f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=rt.SHADE_LAMBERT, shadows=True,
                          reflections=True, plane_y=float(c["plane_y"]))
        col, t, _, _ = s.render(f)
        col2, t2, _, _ = s.render(f, count=False)
        assert util.same_bits(t2, t) and np.array_equal(col2, col), f"{key}: plane + secondary rays, flavours differ"
        assert util.same_bits(t, c["t_full"]), f"{key}: plane + secondary rays, depth"
        if kind == "mesh" and c["scene"] != "cube":
            # a secondary ray starts on the surface: whether a triangle behind it is tested depends on the tree (T6)
            assert (col != c["color_full"]).sum() <= 2e-3 * W * H, f"{key}: plane + secondary rays"
        else:  # SDF scenes, and the cube's brute-force path (mesh-box cull + depth bound), are exact
            assert np.array_equal(col, c["color_full"]), f"{key}: plane + secondary rays"


@pytest.mark.parametrize("name,n_cams", (("example_grid", 150), ("sdf_6", 150), ("sdf_5", 100)))
def test_device_source_on_cpu_vs_oracle_seeded_cameras_sdf(name, n_cams):
    """SDF scenes are tree-free: colour, depth AND the visit counters must equal the oracle's on every camera."""
    s, kind = emul_scene(name)
    ps = util.port_scene(G, name)
    W, H = 72, 48
    pinv = rt.host_projection(W, H)[1]
    for k, pos, vinv in seeded_cameras(n_cams, seed=1000 + len(name)):
        mode = (rt.SHADE_NORMAL, rt.SHADE_LAMBERT, rt.SHADE_COLOR)[k % 3]
        plane = (k % 5 == 0)
        f = rt.make_frame(pos, vinv, pinv, W, H, mode=mode, shadows=plane, reflections=plane,
                          plane_y=-1.0 if plane else None)
        col, t, _, cnt = s.render(f)
        col2, t2, _, _ = s.render(f, count=False)  # the plain flavour the timed launches run
        assert util.same_bits(t2, t) and np.array_equal(col2, col), f"{name} camera {k}: counting and plain flavours differ"
        sc = ps.with_plane(-1.0) if plane else ps
        pc, pt, _, pcnt = sc.render(pos, vinv, pinv, W, H, mode=mode, shadows=plane, reflections=plane, want_counters=True)
        assert util.same_bits(t, pt), f"{name} camera {k}: depth differs on {(util.bits(t) != util.bits(pt)).sum()} pixels"
        assert np.array_equal(col, pc), f"{name} camera {k}: colour differs on {(col != pc).sum()} pixels"
        for cn, pn in COUNTERS[kind].items():
            assert cnt[cn] == pcnt[pn], f"{name} camera {k}: counter {cn} {cnt[cn]} != {pcnt[pn]}"


@pytest.mark.parametrize("name", ("example_grid", "sdf_6", "sdf_5"))
def test_device_source_on_cpu_in_out_depth_buffer(name):
    """raytracing.cpp:90-94: a pixel is replaced only by a nearer hit; the previous depth also bounds the march.  Depths
    in front of tNear and beyond tFar included, eyes inside the box included, with and without the ground plane and the
    secondary rays."""
    s, _ = emul_scene(name)
    ps = util.port_scene(G, name)
    W, H = 64, 40
    pinv = rt.host_projection(W, H)[1]
    rng = np.random.default_rng(5)
    replaced = kept = 0
    for k, pos, vinv in seeded_cameras(60, seed=78):
        t0 = rng.choice(np.array([np.inf, 0.005, 0.05, 0.5, 1.2, 2.0, 2.6, 3.5, 150.0], np.float32), size=(H, W)).astype(np.float32)
        c0 = rng.integers(0, 2**32, size=(H, W), dtype=np.uint32)
        plane = k % 4 == 0
        f = rt.make_frame(pos, vinv, pinv, W, H, mode=rt.SHADE_LAMBERT, shadows=plane, reflections=plane,
                          plane_y=-1.0 if plane else None)
        col, t, _, _ = s.render(f, color=c0, t=t0)
        sc = ps.with_plane(-1.0) if plane else ps
        pc, pt, _, _ = sc.render(pos, vinv, pinv, W, H, mode=P.MODE_LAMBERT, shadows=plane, reflections=plane, color=c0, t=t0)
        assert util.same_bits(t, pt) and np.array_equal(col, pc), f"{name} camera {k}"
        replaced += int((t != t0).sum())
        kept += int((t == t0).sum())
    assert replaced > 0 and kept > 0


@pytest.mark.parametrize("name,brute", (("example_grid", None), ("sdf_6", None), ("sdf_5", None), ("cube", None),
                                        ("cube", True), ("spot", None), ("stanford-bunny", None)))
def test_a_nearer_far_bound_only_rejects(name, brute):
    """What the in/out tiles of render_tiles_staged_kernel rest on ("trace first, fetch later"): the
    previous depth enters draw only as min(tFar, tPrev) (raytracing.cpp:17), so (a) a pixel that is not hit against a
    cleared depth is not hit against ANY previous depth, and (b) a pixel whose previous depth is not below tFar comes out
    exactly as against the cleared one.  The device source on the CPU, eyes inside the box / mesh included, with and
    without the plane and the secondary rays."""
    a = G.scene_arrays(name)
    s = E.EmulScene.mesh(a["vpos4"], a["indices"], brute=brute) if "vpos4" in a else emul_scene(name)[0]
    W, H = 48, 32
    pinv = rt.host_projection(W, H)[1]
    rng = np.random.default_rng(11)
    changed = 0
    for k, pos, vinv in seeded_cameras(16 if name in ("stanford-bunny", "spot") else 40, seed=203):
        plane = k % 3 == 0
        f = rt.make_frame(pos, vinv, pinv, W, H, mode=rt.SHADE_LAMBERT, shadows=plane, reflections=plane,
                          plane_y=-1.0 if plane else None)
        t0 = rng.choice(np.array([np.inf, 0.005, 0.011, 0.3, 0.9, 1.4, 2.0, 2.5, 3.1, 99.9, 100.0, 150.0], np.float32),
                        size=(H, W)).astype(np.float32)
        t0[rng.random((H, W)) < 0.15] = np.float32(np.nan)
        c0 = rng.integers(1, 2**32, size=(H, W), dtype=np.uint32)
        col_c, t_c, prim_c, _ = s.render(f, count=False)                      # cleared: colour 0, depth +inf
        col_p, t_p, prim_p, _ = s.render(f, color=c0, t=t0, count=False)
        hit_c = np.isfinite(t_c)
        touched = (t_p.view(np.uint32) != t0.view(np.uint32)) | (col_p != c0)
        assert not np.any(touched & ~hit_c), f"{name} camera {k}: a pixel without a hit against the cleared depth changed"
        far = ~(t0 < np.float32(100.0))                                       # inf, 100, 150, NaN: min(tFar, tPrev) = tFar
        same = (t_p.view(np.uint32) == t_c.view(np.uint32)) & (col_p == col_c) & (prim_p == prim_c)
        assert np.all(same[far & hit_c]), f"{name} camera {k}: previous depth >= tFar changed a hit"
        assert not np.any(touched[far & ~hit_c])
        changed += int(touched.sum())
    assert changed > 0


@pytest.mark.parametrize("name,n_cams", (("cube", 60), ("stanford-bunny", 12), ("spot", 12), ("as1-oc-214", 12)))
def test_device_source_on_cpu_vs_oracle_seeded_cameras_mesh_outside(name, n_cams):
    """Our own BVH8 (and the brute-force path for the cube) against the oracle's every-triangle loop from outside the
    mesh, where the closest hit does not depend on the tree: depth bit for bit, colour bar exact-t ties."""
    try:
        s, _ = emul_scene(name)
    except Exception:
        pytest.skip(f"{name} is not among the fixtures")
    ps = util.port_scene(G, name)
    W, H = 64, 40
    pinv = rt.host_projection(W, H)[1]
    ties = 0
    for k, pos, vinv in seeded_cameras(n_cams, seed=31 + len(name), radii=(1.8, 2.5, 4.0)):
        col, t, _, _ = s.render(rt.make_frame(pos, vinv, pinv, W, H))
        pc, pt, _, _ = ps.render(pos, vinv, pinv, W, H)
        assert util.same_bits(t, pt), f"{name} camera {k}: depth differs on {(util.bits(t) != util.bits(pt)).sum()} pixels"
        ties += int((col != pc).sum())
    assert ties <= 2e-3 * n_cams * W * H, f"{name}: {ties} colour mismatches"


def test_brute_force_and_bvh_paths_agree_on_the_cube():
    a = G.scene_arrays("cube")
    sb = E.EmulScene.mesh(a["vpos4"], a["indices"], brute=True)
    st = E.EmulScene.mesh(a["vpos4"], a["indices"], brute=False)
    W, H = 64, 40
    pinv = rt.host_projection(W, H)[1]
    for k, pos, vinv in seeded_cameras(40, seed=9, radii=(1.8, 2.5, 4.0)):
        f = rt.make_frame(pos, vinv, pinv, W, H)
        cb, tb, _, _ = sb.render(f)
        ct, tt, _, _ = st.render(f)
        assert util.same_bits(tb, tt), f"camera {k}"
        assert (cb != ct).sum() <= 0.01 * W * H  # shared edges of the cube's faces


@pytest.mark.ref
@pytest.mark.parametrize("name,n_cams", (("cube", 40), ("stanford-bunny", 16)))
def test_device_source_on_cpu_reference_built_tree_inside_and_outside(name, n_cams):
    """rt_scene_create_mesh_bvh8's path (literal slab test + sort8 order over the tree the reference built) against the
    compiled reference itself, eyes inside the mesh included: there unchecked triangle hits make the answer
    tree-dependent, and with the same tree it must still be the reference's, bit for bit, prim ids included."""
    from oracle import ref_oracle as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    a = G.scene_arrays(name)
    ro = R.RefOracle()
    ref = ro.mesh(a["vpos4"], a["indices"])
    m = ref.export_mesh()
    s = E.EmulScene.mesh_refbvh(m["vpos4"], m["indices"], m["nodes"], m["prim_of_permuted"])
    W, H = 56, 40
    pinv = ro.projection(W, H)[1]
    rng = np.random.default_rng(123)
    for k in range(n_cams):
        v = rng.normal(size=3)
        eye = tuple(float(x) for x in float(rng.choice([0.2, 0.5, 0.9, 1.6, 2.5])) * v / np.linalg.norm(v))
        cam = R.CamSpec(eye=eye, target=tuple(float(x) for x in rng.uniform(-0.3, 0.3, size=3)), lock_up=bool(k % 2),
                        n_rotate=int(rng.integers(0, 4)), dx=float(rng.uniform(-80, 80)), dy=float(rng.uniform(-40, 40)))
        pos, _view, vinv = ro.camera(cam)
        for mode_ref, mode in ((R.MODE_NORMAL, rt.SHADE_NORMAL), (R.MODE_LAMBERT, rt.SHADE_LAMBERT)):
            want_c, want_t, _ = ref.render(cam, W, H, proj_inv=pinv, mode=mode_ref)
            col, t, _, _ = s.render(rt.make_frame(pos, vinv, pinv, W, H, mode=mode))
            assert util.same_bits(t, want_t), f"{name} camera {k} {cam}: depth differs"
            assert np.array_equal(col, want_c), f"{name} camera {k} {cam}: colour differs"


def test_brick_layout_is_the_linear_layout():
    """RT_GRID_BRICKS only changes where a sample lives: frames and sample counts are those of the linear layout."""
    a = G.scene_arrays("example_grid")
    lin, bri = E.EmulScene.grid(a["values"]), E.EmulScene.grid(a["values"], bricks=True)
    W, H = 72, 48
    pinv = rt.host_projection(W, H)[1]
    for k, pos, vinv in seeded_cameras(60, seed=4):
        f = rt.make_frame(pos, vinv, pinv, W, H, mode=rt.SHADE_LAMBERT)
        c0, t0, _, n0 = lin.render(f)
        c1, t1, _, n1 = bri.render(f)
        assert util.same_bits(t0, t1) and np.array_equal(c0, c1) and n0 == n1, f"camera {k}"
    # a grid whose size is not a multiple of 8 in any axis: partial bricks
    v = np.ascontiguousarray(a["values"][:61, 3:62, :59])
    lin, bri = E.EmulScene.grid(v), E.EmulScene.grid(v, bricks=True)
    for k, pos, vinv in seeded_cameras(20, seed=5):
        f = rt.make_frame(pos, vinv, pinv, W, H)
        c0, t0, _, _ = lin.render(f)
        c1, t1, _, _ = bri.render(f)
        assert util.same_bits(t0, t1) and np.array_equal(c0, c1), f"partial bricks, camera {k}"


def _ray_batch(seed, n_random=1500):
    """Exactly axis-parallel rays (a zero direction component makes 1/d infinite and 0 * inf = NaN in the slab tests,
    SURVEY T11) from origins on cell boundaries, box faces and box corners, plus random rays from inside and outside."""
    rng = np.random.default_rng(seed)
    o, d = [], []
    axes = np.eye(3, dtype=np.float32)
    lattice = np.array([-1.0, -0.5, -1.0 / 32, 0.0, 1.0 / 64, 0.25, 1.0], np.float32)
    for ax in range(3):
        for sgn in (1.0, -1.0):
            for u in lattice:
                for v in lattice:
                    p = np.zeros(3, np.float32)
                    p[ax] = -2.5 * sgn
                    p[(ax + 1) % 3], p[(ax + 2) % 3] = u, v
                    o.append(p)
                    d.append(axes[ax] * np.float32(sgn))
                    q = p.copy()
                    q[ax] = np.float32(0.125 * sgn)  # starting inside the box
                    o.append(q)
                    d.append(axes[ax] * np.float32(sgn))
    # two zero components / one zero component
    for _ in range(200):
        p = rng.uniform(-1.5, 1.5, 3).astype(np.float32)
        v = rng.normal(size=3).astype(np.float32)
        v[rng.integers(0, 3)] = 0.0
        o.append(p)
        d.append(v / np.linalg.norm(v))
    for _ in range(n_random):
        p = (rng.uniform(-1, 1, 3) * rng.choice([0.5, 1.0, 3.0])).astype(np.float32)
        v = rng.normal(size=3).astype(np.float32)
        o.append(p)
        d.append((v / np.linalg.norm(v)).astype(np.float32))
    return np.asarray(o, np.float32), np.asarray(d, np.float32)


@pytest.mark.parametrize("name", ("example_grid", "sdf_6", "sdf_5", "cube"))
def test_device_source_on_cpu_single_rays_vs_oracle(name):
    """IScene::intersect (rays_kernel's body): hit flag, t and normal bit for bit, axis-parallel rays included.
    (cube: rays that start inside a mesh are tree-dependent; the explicit RT_MESH_BRUTE rule is the oracle loop's.)"""
    s, kind = emul_scene(name, **({"brute": True} if name == "cube" else {}))
    ps = util.port_scene(G, name)
    o, d = _ray_batch(seed=len(name))
    for t_near, t_far, plane in ((0.01, 100.0, None), (0.3, 2.2, None), (0.01, 100.0, -1.0)):
        hit, t, nrm, _ = s.intersect(o, d, t_near, t_far, plane_y=plane)
        sc = ps.with_plane(plane) if plane is not None else ps
        for i in range(o.shape[0]):
            h, pt, pn, _ = sc.intersect(o[i], d[i], t_near, t_far)
            assert bool(hit[i]) == h, f"{name} ray {i} o={o[i]} d={d[i]}: hit {hit[i]} vs {h}"
            if h:
                assert util.same_bits(t[i], pt), f"{name} ray {i} o={o[i]} d={d[i]}: t {t[i]} vs {pt}"
                assert util.same_bits(nrm[i], pn), f"{name} ray {i} o={o[i]} d={d[i]}: normal {nrm[i]} vs {pn}"


@pytest.mark.ref
def test_brute_force_path_matches_the_reference():
    """The <= 32-triangle kernel has no tree; the reference reaches triangles only through boxes that pass its slab test.
    The mesh-bounding-box cull plus `t <= tFar` per triangle reproduce it on the cube (BASELINE configs[0]): frames from
    outside, from just outside the faces and from inside, looking at the cube and away from it, on cleared and
    pre-filled depth buffers, and single rays with arbitrary [tNear, tFar].  Both the kernel source and the oracle's
    every-triangle loop."""
    from oracle import ref_oracle as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    a = G.scene_arrays("cube")
    ro = R.RefOracle()
    ref = ro.mesh(a["vpos4"], a["indices"])
    s = E.EmulScene.mesh(a["vpos4"], a["indices"], brute=True)
    ps = P.Scene.mesh_brute(a["vpos4"], a["indices"])
    W, H = 64, 40
    pinv = ro.projection(W, H)[1]
    rng = np.random.default_rng(3)
    for radii in ((1.8, 2.5, 4.0), (1.05, 1.2), (0.2, 0.5, 0.9)):
        for prefill in (False, True):
            for away in (False, True):
                for k in range(12):
                    v = rng.normal(size=3)
                    v /= np.linalg.norm(v)
                    r = float(rng.choice(radii))
                    eye = tuple(float(x) for x in r * v)
                    target = (tuple(float(x) for x in 2.5 * r * v + rng.uniform(-0.5, 0.5, 3)) if away
                              else tuple(float(x) for x in rng.uniform(-0.3, 0.3, size=3)))
                    cam = R.CamSpec(eye=eye, target=target, lock_up=bool(k % 2))
                    pos, _view, vinv = ro.camera(cam)
                    t0 = (rng.choice(np.array([np.inf, 0.5, 1.2, 2.0, 2.6, 3.5], np.float32), size=(H, W)).astype(np.float32)
                          if prefill else np.full((H, W), np.inf, np.float32))
                    c0 = rng.integers(0, 2**32, size=(H, W), dtype=np.uint32)
                    want_c, want_t, _ = ref.render(cam, W, H, proj_inv=pinv, color=c0, t=t0)
                    col, t, _, _ = s.render(rt.make_frame(pos, vinv, pinv, W, H), color=c0, t=t0)
                    pc, pt, _, _ = ps.render(pos, vinv, pinv, W, H, color=c0, t=t0)
                    where = f"radii {radii} prefill {prefill} away {away} camera {k}"
                    assert util.same_bits(t, want_t), f"kernel source, {where}: depth differs"
                    assert util.same_bits(pt, want_t), f"oracle, {where}: depth differs"
                    assert (col != want_c).sum() <= 0.01 * W * H and (pc != want_c).sum() <= 0.01 * W * H  # edge ties
    o, d = _ray_batch(seed=4, n_random=800)
    for t_near, t_far in ((0.01, 100.0), (0.3, 2.2), (1.0, 1.6)):
        hit, t, _, _ = s.intersect(o, d, t_near, t_far)
        for i in range(o.shape[0]):
            r = ref.intersect(o[i], d[i], t_near, t_far)
            h, pt, _, _ = ps.intersect(o[i], d[i], t_near, t_far)
            assert bool(hit[i]) == bool(r[0]) == h, f"ray {i} o={o[i]} d={d[i]} [{t_near}, {t_far}]"
            if h:
                assert util.same_bits(t[i], r[1]) and util.same_bits(pt, r[1]), f"ray {i} [{t_near}, {t_far}]"


def test_brute_force_path_golden_prefilled_frames():
    a = G.scene_arrays("cube")
    s = E.EmulScene.mesh(a["vpos4"], a["indices"], brute=True)
    for key in [c for c in G.cases if c.startswith("cube|")]:
        c = G.case(key)
        t0 = np.full((c["H"], c["W"]), 2.2, np.float32)
        c0 = np.full((c["H"], c["W"]), 0x11223344, np.uint32)
        col, t, _, _ = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"]), color=c0, t=t0)
        assert util.same_bits(t, c["t_prefilled"]) and np.array_equal(col, c["color_prefilled"]), key


def test_device_source_on_cpu_large_standins():
    """BASELINE configs[2] / [3] stand-ins (257^3 grid in both layouts, depth-8 octree with 0.6 M nodes) on seeded
    cameras, visit counters included: the deeper tree and the larger index arithmetic, which the GPU tests only see
    from the default camera."""
    W, H = 72, 48
    pinv = rt.host_projection(W, H)[1]
    a = G.scene_arrays("octree_d8")
    s, ps = E.EmulScene.octree(a["nodes36"]), P.Scene.octree(a["nodes36"])
    for k, pos, vinv in seeded_cameras(40, seed=12):
        col, t, _, cnt = s.render(rt.make_frame(pos, vinv, pinv, W, H, mode=rt.SHADE_LAMBERT))
        pc, pt, _, pcnt = ps.render(pos, vinv, pinv, W, H, mode=P.MODE_LAMBERT, want_counters=True)
        assert util.same_bits(t, pt) and np.array_equal(col, pc), f"octree_d8 camera {k}"
        for cn, pn in COUNTERS["octree"].items():
            assert cnt[cn] == pcnt[pn], f"octree_d8 camera {k}: counter {cn}"
    a = G.scene_arrays("example_grid_x4")
    lin, bri, ps = E.EmulScene.grid(a["values"]), E.EmulScene.grid(a["values"], bricks=True), P.Scene.grid(a["values"])
    for k, pos, vinv in seeded_cameras(30, seed=13):
        f = rt.make_frame(pos, vinv, pinv, W, H)
        pc, pt, _, pcnt = ps.render(pos, vinv, pinv, W, H, want_counters=True)
        for s in (lin, bri):
            col, t, _, cnt = s.render(f)
            assert util.same_bits(t, pt) and np.array_equal(col, pc), f"grid x4 camera {k}"
            assert cnt["grid_samples"] == pcnt["gridSdf"]


@pytest.mark.parametrize("name", ("sdf_6", "sdf_5", "octree_d8"))
def test_resumable_octree_walk_equals_the_plain_traversal(name):
    """OctWalk (begin / step / leaf: the lane state machine of render_octree_compact_kernel) driven to completion for one
    ray at a time visits the same nodes, marches the same leaves and returns the same hits as oct_intersect."""
    a = G.scene_arrays(name)
    s = E.EmulScene.octree(a["nodes36"])
    o, d = _ray_batch(seed=21, n_random=4000)
    for t_near, t_far in ((0.01, 100.0), (0.3, 2.2)):
        h0, t0, n0, c0 = s.octree_rays(o, d, t_near, t_far, walk=False)
        h1, t1, n1, c1 = s.octree_rays(o, d, t_near, t_far, walk=True)
        assert np.array_equal(h0, h1) and util.same_bits(t0, t1) and util.same_bits(n0[h0], n1[h0])
        assert c0 == c1, f"{name}: visit counters {c0} vs {c1}"


def _random_octree(rng, max_depth, p_inner):
    """A structurally valid random octree in the file's 36-byte records (8 children contiguous at childrenOffset), with
    leaf values drawn from a mix that hits every classification rule: empty markers (all 1000 / all 0), all-positive
    leaves just above and below the 1e-4 threshold, sign changes, exact zeros at corners, large magnitudes."""
    vals, offs = [np.zeros(8, np.float32)], [0]
    frontier = [(0, 0)]
    while frontier:
        nxt = []
        for node, depth in frontier:
            if depth < max_depth and (depth == 0 or rng.random() < p_inner):
                first = len(vals)
                offs[node] = first
                for _ in range(8):
                    vals.append(np.zeros(8, np.float32))
                    offs.append(0)
                    nxt.append((len(vals) - 1, depth + 1))
            else:
                kind = rng.integers(0, 8)
                if kind == 0:
                    v = np.full(8, 1000.0)
                elif kind == 1:
                    v = np.zeros(8)
                elif kind == 2:
                    v = rng.uniform(1e-4, 3e-4, 8)          # all >= 1e-4: rejected without a march
                elif kind == 3:
                    v = rng.uniform(0.5e-4, 2e-4, 8)        # straddles the threshold
                elif kind == 4:
                    v = rng.uniform(-0.2, 0.3, 8)
                    v[rng.integers(0, 8)] = 0.0             # an exact zero at a corner
                elif kind == 5:
                    v = rng.uniform(-0.05, 0.6, 8)
                elif kind == 6:
                    v = rng.uniform(5.0, 20.0, 8)           # > 10 everywhere or mixed around it
                else:
                    v = rng.normal(0.1, 0.2, 8)
                vals[node] = v.astype(np.float32)
        frontier = nxt
    rec = np.zeros((len(vals), 36), np.uint8)
    rec[:, :32] = np.asarray(vals, np.float32).view(np.uint8).reshape(-1, 32)
    rec[:, 32:] = np.asarray(offs, np.uint32).view(np.uint8).reshape(-1, 4)
    return rec


@pytest.mark.parametrize("seed", range(12))
def test_random_octrees_vs_oracle(seed):
    """Structures the shipped files never contain (sparse and lopsided trees, leaves at every depth, every leaf class
    next to every other): frames with counters, single rays incl. axis-parallel ones, and the resumable walk."""
    rng = np.random.default_rng(900 + seed)
    nodes = _random_octree(rng, max_depth=int(rng.integers(1, 5)), p_inner=float(rng.uniform(0.2, 0.7)))
    s, ps = E.EmulScene.octree(nodes), P.Scene.octree(nodes)
    W, H = 48, 32
    pinv = rt.host_projection(W, H)[1]
    for k, pos, vinv in seeded_cameras(10, seed=seed):
        col, t, _, cnt = s.render(rt.make_frame(pos, vinv, pinv, W, H))
        col2, t2, _, _ = s.render(rt.make_frame(pos, vinv, pinv, W, H), count=False)
        assert util.same_bits(t2, t) and np.array_equal(col2, col), "counting and plain flavours differ"
        pc, pt, _, pcnt = ps.render(pos, vinv, pinv, W, H, want_counters=True)
        assert util.same_bits(t, pt) and np.array_equal(col, pc), f"seed {seed} camera {k} ({nodes.shape[0]} nodes)"
        for cn, pn in COUNTERS["octree"].items():
            assert cnt[cn] == pcnt[pn], f"seed {seed} camera {k}: counter {cn} {cnt[cn]} != {pcnt[pn]}"
    o, d = _ray_batch(seed=seed, n_random=300)
    hit, t, nrm, _ = s.intersect(o, d)
    for i in range(o.shape[0]):
        h, pt, pn, _ = ps.intersect(o[i], d[i])
        assert bool(hit[i]) == h and (not h or (util.same_bits(t[i], pt) and util.same_bits(nrm[i], pn))), f"seed {seed} ray {i}"
    h1, t1, n1, c1 = s.octree_rays(o, d, walk=True)
    h0, t0, n0, c0 = s.octree_rays(o, d, walk=False)
    assert np.array_equal(h0, h1) and util.same_bits(t0, t1) and c0 == c1


@pytest.mark.parametrize("seed", range(10))
def test_random_grids_vs_oracle(seed):
    """Grids the shipped file is not: 2 samples on an axis, non-cubic sizes, sizes that are not a multiple of the brick
    edge, random (non-distance) fields with the surface touching the box faces; both layouts."""
    rng = np.random.default_rng(500 + seed)
    size = tuple(int(x) for x in rng.integers(2, 12, size=3))
    x, y, z = np.meshgrid(*(np.linspace(-1, 1, n) for n in size), indexing="ij")
    c = rng.uniform(-0.5, 0.5, 3)
    field = np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - rng.uniform(0.3, 1.2)
    field += rng.normal(0, 0.05, size)
    v = np.ascontiguousarray(field, np.float32)
    lin, bri, ps = E.EmulScene.grid(v), E.EmulScene.grid(v, bricks=True), P.Scene.grid(v)
    W, H = 48, 32
    pinv = rt.host_projection(W, H)[1]
    for k, pos, vinv in seeded_cameras(10, seed=seed):
        pc, pt, _, pcnt = ps.render(pos, vinv, pinv, W, H, mode=P.MODE_LAMBERT, want_counters=True)
        for s in (lin, bri):
            col, t, _, cnt = s.render(rt.make_frame(pos, vinv, pinv, W, H, mode=rt.SHADE_LAMBERT))
            assert util.same_bits(t, pt) and np.array_equal(col, pc), f"seed {seed} size {size} camera {k}"
            assert cnt["grid_samples"] == pcnt["gridSdf"]
    o, d = _ray_batch(seed=seed, n_random=300)
    hit, t, nrm, _ = lin.intersect(o, d)
    for i in range(o.shape[0]):
        h, pt, pn, _ = ps.intersect(o[i], d[i])
        assert bool(hit[i]) == h and (not h or (util.same_bits(t[i], pt) and util.same_bits(nrm[i], pn))), f"seed {seed} ray {i}"


def test_one_node_octrees_vs_oracle():
    """A file whose root is a leaf (OctreeDev::root_inner = 0): every leaf class."""
    W, H = 48, 32
    pinv = rt.host_projection(W, H)[1]
    rng = np.random.default_rng(1)
    leaves = (np.full(8, 1000.0), np.zeros(8), rng.uniform(-0.5, 0.5, 8), rng.uniform(-0.2, 0.9, 8),
              rng.uniform(1e-4, 2e-4, 8), np.array([0.5, 0.5, 0.5, 0.5, -0.5, -0.5, -0.5, -0.5]))
    some_hit = False
    for case, v in enumerate(leaves):
        rec = np.zeros((1, 36), np.uint8)
        rec[0, :32] = v.astype(np.float32).view(np.uint8)
        s, ps = E.EmulScene.octree(rec), P.Scene.octree(rec)
        for k, pos, vinv in seeded_cameras(20, seed=case):
            col, t, _, cnt = s.render(rt.make_frame(pos, vinv, pinv, W, H))
            pc, pt, _, pcnt = ps.render(pos, vinv, pinv, W, H, want_counters=True)
            assert util.same_bits(t, pt) and np.array_equal(col, pc), f"leaf {case} camera {k}"
            assert cnt["oct_leaf_samples"] == pcnt["octLeafSdf"]
            some_hit |= bool(np.isfinite(t).any())
    assert some_hit
