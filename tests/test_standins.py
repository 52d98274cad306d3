"""The deterministic stand-ins for the benchmark resources the reference tree does not ship:
format conventions and invariants on CPU; the CUDA path against the oracle at BASELINE.json's full sizes on GPU."""
import numpy as np
import pytest

from tests import util
from triangles_sdf_cpu_raytracing_b200 import standins as S

G = util.Golden()
ORBIT_DX = float(np.float32(2.0 * np.pi / 64.0 / 0.01))  # one 64th of a turn per Camera::rotate (camera.hpp:55)


# ------------------------------------------------------------------------------------------------ CPU: generators
def test_refined_grid_keeps_the_coarse_samples_and_the_picture():
    from oracle import port_oracle as P
    v = G.scene_arrays("example_grid")["values"]
    v4 = S.refine_grid(v, 4)
    assert v4.shape == (257, 257, 257) and v4.dtype == np.float32
    assert np.array_equal(v4[::4, ::4, ::4], v)
    # refinement of a trilinear interpolant is the same function: along a z edge the new samples are linear
    assert np.allclose(v4[0, 0, 2], 0.5 * (v[0, 0, 0].astype(np.float64) + v[0, 0, 1]), rtol=0, atol=1e-7)
    c = G.case("example_grid|default|160x90")
    a = P.Scene.grid(v).render(c["rayPos"], c["viewInv"], c["projInv"], 160, 90)
    b = P.Scene.grid(v4).render(c["rayPos"], c["viewInv"], c["projInv"], 160, 90)
    d = util.compare_frames(b[0], b[1], a[0], a[1])
    assert d["mask_mismatch"] == 0 and d["depth_outliers"] == 0 and d["rgb_outliers"] <= 2, d
    assert len(S.grid_file_bytes(v4)) == 12 + 4 * 257 ** 3


def test_octree_from_grid_follows_the_shipped_conventions():
    v = G.scene_arrays("example_grid")["values"]
    rec = S.octree_from_grid(v, 6)
    vals = rec[:, :32].copy().view(np.float32).reshape(-1, 8)
    off = rec[:, 32:].copy().view(np.uint32).reshape(-1)
    n = rec.shape[0]
    inner = off != 0
    assert off[0] == 1 and n == 1 + 8 * int(inner.sum())          # root's children at 1..8, blocks of 8, no holes
    assert np.all(vals[inner] == 0.0)                               # inner nodes carry all-zero values
    assert np.all(off[inner] + 7 < n) and np.all((off[inner] - 1) % 8 == 0)
    assert np.array_equal(np.sort(off[inner]), 1 + 8 * np.arange(int(inner.sum())))  # each block has one parent
    leaves = vals[~inner]
    empty_out, empty_in = np.all(leaves == 1000.0, axis=1), np.all(leaves == 0.0, axis=1)
    surface = ~(empty_out | empty_in)
    assert surface.sum() > 1000 and np.all(leaves[surface].min(axis=1) < 0) and np.all(leaves[surface].max(axis=1) > 0)
    # same order of magnitude as the shipped depth-6 octree (36 417 nodes)
    assert 20000 < n < 60000
    # the port oracle renders it like the grid it came from (different HIT_EPS: depth within 1e-2, mask bar edges)
    from oracle import port_oracle as P
    c = G.case("example_grid|default|160x90")
    a = P.Scene.grid(v).render(c["rayPos"], c["viewInv"], c["projInv"], 160, 90)
    b = P.Scene.octree(rec).render(c["rayPos"], c["viewInv"], c["projInv"], 160, 90)
    d = util.compare_frames(b[0], b[1], a[0], a[1], depth_rtol=1e-2)
    assert d["mask_mismatch"] <= 0.01 * 160 * 90 and d["depth_outliers"] <= 0.01 * 160 * 90, d
    assert len(S.octree_file_bytes(rec)) == 4 + 36 * n


def test_subdivided_mesh_is_the_same_surface():
    m = G.scene_arrays("cube")
    v, i = S.subdivide_mesh(m["vpos4"], m["indices"], 2)
    assert i.size == 16 * m["indices"].size and np.all(v[:, 3] == 1.0)
    def area(v, i):
        p = v[i.reshape(-1, 3), :3].astype(np.float64)
        return 0.5 * np.linalg.norm(np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]), axis=1)
    a0, a1 = area(m["vpos4"], m["indices"]), area(v, i)
    assert np.allclose(a1.reshape(-1, 16).sum(axis=1), a0, rtol=1e-5)   # children 16k..16k+15 tile parent k
    from oracle import port_oracle as P
    c = G.case("cube|default|128x128")
    a = P.Scene.mesh_brute(m["vpos4"], m["indices"]).render(c["rayPos"], c["viewInv"], c["projInv"], 64, 64)
    b = P.Scene.mesh_brute(v, i).render(c["rayPos"], c["viewInv"], c["projInv"], 64, 64)
    d = util.compare_frames(b[0], b[1], a[0], a[1])
    assert d["mask_mismatch"] <= 8 and d["depth_outliers"] == 0, d
    both = np.isfinite(a[1]) & np.isfinite(b[1])
    # child prim ids map back to their parent (bar pixels on an edge shared by two triangles, where t ties)
    assert np.count_nonzero(b[2][both] // 16 != a[2][both]) <= 0.02 * both.sum()


# ------------------------------------------------------------------------------------------------ GPU: full sizes
def _cam(W, H):
    c = G.case("example_grid|default|160x90")
    assert abs(W / H - 16 / 9) < 1e-6
    return c


@pytest.mark.gpu
@pytest.mark.parametrize("name,W,H", (("example_grid_x4", 3840, 2160), ("example_grid_x8", 3840, 2160),
                                      ("octree_d8", 3840, 2160), ("octree_d9", 3840, 2160)))
def test_large_sdf_standins_bit_exact_at_full_size(name, W, H):
    """BASELINE.json configs[2], [3]: the large grid / octree stand-ins at 3840x2160 against the plain-C oracle."""
    import triangles_sdf_cpu_raytracing_b200 as rt
    rt.set_math_mode(rt.MATH_EXACT)
    c = _cam(W, H)
    want_c, want_t, want_p, want_cnt = util.port_scene(G, name).render(c["rayPos"], c["viewInv"], c["projInv"], W, H,
                                                                       want_counters=True)
    layouts = (rt.GRID_LINEAR, rt.GRID_BRICKS) if name.startswith("example_grid") else (None,)
    for layout in layouts:
        s = util.cuda_scene(G, name, **({"layout": layout} if layout is not None else {}))
        for sched in (rt.SCHED_TILES, rt.SCHED_PERSISTENT, rt.SCHED_COMPACT):
            rt.set_sched_mode(sched)
            r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), want_prim=True, want_counters=True)
            assert util.same_bits(r["t"], want_t) and np.array_equal(r["color"], want_c), (name, layout, sched)
            assert np.array_equal(r["prim"], want_p)
            cn = r["counters"]
            assert cn["grid_samples"] == want_cnt["gridSdf"] and cn["oct_inner"] == want_cnt["octInner"]
            assert cn["oct_leaf"] == want_cnt["octLeaf"] and cn["oct_leaf_samples"] == want_cnt["octLeafSdf"]
        s.destroy()
    rt.set_sched_mode(rt.SCHED_AUTO)


@pytest.mark.gpu
@pytest.mark.ref
def test_large_mesh_standin_orbit_vs_reference():
    """BASELINE.json configs[4]: the x16-subdivided bunny (1.1 M triangles) at 3840x2160 on orbit cameras, against
    the reference's own BVHBuilder + Renderer::draw (oracle/_ref)."""
    from oracle import port_oracle as P
    from oracle import ref_oracle as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    import triangles_sdf_cpu_raytracing_b200 as rt
    rt.set_math_mode(rt.MATH_EXACT)
    a = G.scene_arrays("stanford-bunny_s2")
    assert a["indices"].size // 3 == 16 * 69451
    W, H = 3840, 2160
    c = _cam(W, H)
    ref = R.RefOracle().mesh(a["vpos4"], a["indices"])
    s = rt.TriangleMesh(a["vpos4"], a["indices"])
    for k in (0, 8, 17, 24, 32, 40, 48, 56, 63):  # frames of the 64-frame orbit
        pos, _view, vinv = rt.host_camera(n_rotate=k, dx=ORBIT_DX, dy=0.0)
        want_c, want_t, _ = ref.render(R.CamSpec(n_rotate=k, dx=ORBIT_DX, dy=0.0), W, H, proj_inv=c["projInv"])
        r = s.render(rt.make_frame(pos, vinv, c["projInv"], W, H), want_prim=True)
        assert util.same_bits(r["t"], want_t), f"orbit frame {k}: depth differs"
        # Colour may differ only where two triangles give the bit-identical closest t (a shared edge): the winner, and
        # with it the normal, then depends on the visit order, i.e. on the tree.  Each such pixel is
        # re-checked with the oracle's single-triangle test on OUR winner; they are counted and bounded.
        ys, xs = np.nonzero(r["color"] != want_c)
        assert ys.size <= 1e-4 * W * H, f"orbit frame {k}: {ys.size} colour mismatches"
        v, idx = a["vpos4"].reshape(-1, 4), a["indices"].reshape(-1, 3)
        for row, x in zip(ys, xs):
            d = P.eye_ray(pos, vinv, c["projInv"], W, H, int(x), int(H - 1 - row))
            tri = v[idx[r["prim"][row, x]], :3] / v[idx[r["prim"][row, x]], 3:4]
            hit, tt, _ = P.triangle(pos, d, tri.astype(np.float32))
            assert hit and util.same_bits(tt, want_t[row, x]), f"orbit frame {k} pixel ({x},{row}) is not a tie"
    s.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("name,key,W,H", (("cube", "cube|default|128x128", 512, 512),
                                          ("example_grid", "example_grid|default|160x90", 1920, 1080)))
def test_baseline_configs_0_and_1_at_their_own_size(name, key, W, H):
    """BASELINE.json configs[0] (cube.obj 512x512) and configs[1] (example_grid.grid 1920x1080) at exactly those sizes:
    against the reference's own Renderer::draw where oracle/_ref travelled to the box, and against the plain-C oracle."""
    import triangles_sdf_cpu_raytracing_b200 as rt
    from oracle import ref_oracle as R
    rt.set_math_mode(rt.MATH_EXACT)
    rt.set_sched_mode(rt.SCHED_AUTO)
    c = G.case(key)
    assert abs(W / H - c["W"] / c["H"]) < 1e-6
    a = G.scene_arrays(name)
    wants = []
    pc, pt, _, _ = util.port_scene(G, name).render(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    wants.append(("oracle/port_oracle.c", pc, pt))
    if R.available():
        o = R.RefOracle()
        ref = o.mesh(a["vpos4"], a["indices"]) if "vpos4" in a else o.grid(a["values"])
        rc, rtt, _ = ref.render(R.CamSpec(), W, H, proj_inv=c["projInv"])
        wants.append(("oracle/_ref", rc, rtt))
    accels = (rt.MESH_AUTO, rt.MESH_BRUTE, rt.MESH_BVH8) if name == "cube" else (None,)
    for accel in accels:
        s = util.cuda_scene(G, name, **({"accel": accel} if accel is not None else {}))
        r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H))
        rp = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True),
                      color=np.full((H, W), 0x5A5A5A5A, np.uint32), t=np.full((H, W), 0.25, np.float32))
        for who, wc, wt in wants:
            assert util.same_bits(r["t"], wt), f"{name} {W}x{H} accel {accel}: depth differs from {who}"
            assert util.same_bits(rp["t"], wt), f"{name} {W}x{H} accel {accel}, fused clear: depth differs from {who}"
            if name == "cube":  # the two triangles of a face tie exactly along their shared diagonal: same normal
                assert np.array_equal(r["color"], wc) and np.array_equal(rp["color"], wc), (who, accel)
            else:
                assert np.array_equal(r["color"], wc) and np.array_equal(rp["color"], wc), who
        s.destroy()
