"""GPU: one process, several devices (rt_init, include/rt_b200.h) and the extended kernel-stores-to-host path.

The reference renders a frame with ONE call from ONE process (src/main.cpp:203,206 -> Renderer::draw,
src/raytracing.cpp:67-102).  After rt_init(devs, n) the same single rt_render_host / rt_render_device call fans the
frame out over n GPUs; everything that comes back must still be the reference's frame, bit for bit.
rt_init is process-wide, so each scenario runs in its own interpreter.  The multi-device scenarios skip on a 1-GPU box
(`gpurun --gpus 2` runs them); the single-device ones always run."""
from __future__ import annotations

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRELUDE = """
import ctypes as C
import numpy as np, torch
import triangles_sdf_cpu_raytracing_b200 as rt
from triangles_sdf_cpu_raytracing_b200 import _lib as L
from tests import util
g = util.Golden()
rt.set_math_mode(rt.MATH_EXACT)

def pinned(H, W, dtype, fill):
    t = torch.full((H * W,), fill, dtype=dtype).pin_memory()
    return t

def host_call(s, f, color, t, prim=None, cnt=None):
    dms, tms = C.c_float(), C.c_float()
    L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(color.data_ptr()), C.c_void_p(t.data_ptr()),
                                   C.c_void_p(prim.data_ptr()) if prim is not None else None, C.byref(dms), C.byref(tms),
                                   C.byref(cnt) if cnt is not None else None))
    return dms.value, tms.value

def full_frame(c, **kw):
    return rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], mode=rt.SHADE_LAMBERT, shadows=True,
                         reflections=True, plane_y=float(c["plane_y"]), **kw)

def check_scene(key, s, tag):
    c = g.case(key)
    W, H = c["W"], c["H"]
    is_mesh = c["scene"] in util.MESHES
    # --- page-locked buffers, fused clear over dirty contents, then the in/out contract
    color, t = pinned(H, W, torch.int32, 0x5A5A5A5A), pinned(H, W, torch.float32, 0.25)
    cn, tn = color.numpy().view(np.uint32).reshape(H, W), t.numpy().reshape(H, W)
    host_call(s, rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True), color, t)
    assert util.same_bits(tn, c["t"]) and np.array_equal(cn, c["color_normal"]), (tag, key, "pinned clear")
    cn[:, :W // 2] = 0; tn[:, :W // 2] = np.inf; cn[:, W // 2:] = 7; tn[:, W // 2:] = 1e-3
    host_call(s, rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), color, t)
    assert util.same_bits(tn[:, :W // 2], c["t"][:, :W // 2]) and np.array_equal(cn[:, :W // 2], c["color_normal"][:, :W // 2]), (tag, key, "pinned in/out")
    assert np.all(cn[:, W // 2:] == 7) and np.all(tn[:, W // 2:] == np.float32(1e-3)), (tag, key, "nearer depth survives")
    # --- pinned in/out over a pre-filled depth (golden), with the transfer accounting of that one call: stored = the
    #     tiles (32 x 4 pixels) that changed, fetched = at least their depth and at most every tile's depth + the colour
    #     of the partly changed ones (tiles no ray can reach the model's bounds in are not even read)
    cn[...] = 0x11223344; tn[...] = 2.2
    f0, s0 = s.transfer_bytes()
    host_call(s, rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), color, t)
    f1, s1 = s.transfer_bytes()
    if not is_mesh or c["scene"] == "cube":
        assert util.same_bits(tn, c["t_prefilled"]) and np.array_equal(cn, c["color_prefilled"]), (tag, key, "pinned prefilled")
    if W % 32 == 0:
        changed = tn != np.float32(2.2)
        stored = fetched_min = fetched_max = 0
        for y0 in range(0, H, 4):
            rows = min(4, H - y0)
            for x0 in range(0, W, 32):
                n = int(changed[y0:y0 + rows, x0:x0 + 32].sum())
                stored += rows * 32 * 8 if n else 0
                fetched_min += rows * 32 * 4 * ((1 if n else 0) + (1 if 0 < n < rows * 32 else 0))
                fetched_max += rows * 32 * 4 * (1 + (1 if 0 < n < rows * 32 else 0))
        assert s1 - s0 == stored and fetched_min <= f1 - f0 <= fetched_max, (tag, key, "transfer bytes", f1 - f0, s1 - s0, stored, fetched_min, fetched_max)
        if c["scene"] == "cube":
            assert f1 - f0 < fetched_max, (tag, key, "tiles off the model are not fetched", f1 - f0, fetched_max)
    # --- pinned prim ids: equal to what the copy pipeline returns (counters force the copies)
    prim = pinned(H, W, torch.int32, 0x33333333)
    host_call(s, rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True), color, t, prim=prim)
    r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), want_prim=True, want_counters=True)
    assert util.same_bits(tn, r["t"]) and np.array_equal(prim.numpy().view(np.uint32).reshape(H, W), r["prim"]), (tag, key, "prim")
    assert util.same_bits(r["t"], c["t"]) and np.array_equal(r["color"], c["color_normal"]), (tag, key, "pageable")
    assert r["counters"]["rays"] == W * H, (tag, key, r["counters"])
    # --- pageable, pre-filled depth (golden), Lambert
    t0 = np.full((H, W), 2.2, np.float32); c0 = np.full((H, W), 0x11223344, np.uint32)
    r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), color=c0, t=t0)
    if not is_mesh or c["scene"] == "cube":
        assert util.same_bits(r["t"], c["t_prefilled"]) and np.array_equal(r["color"], c["color_prefilled"]), (tag, key, "prefilled")
    r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=rt.SHADE_LAMBERT))
    assert np.array_equal(r["color"], c["color_lambert"]), (tag, key, "lambert")
    # --- the viewer's default look (plane + shadows + reflections) through pinned and pageable buffers
    if c["color_full"] is not None:
        host_call(s, full_frame(c, clear=True), color, t)
        r = s.render(full_frame(c))
        assert util.same_bits(tn, r["t"]) and np.array_equal(cn, r["color"]), (tag, key, "default look, pinned vs pageable")
        assert util.same_bits(tn, c["t_full"]), (tag, key, "default look depth")
        if not is_mesh or c["scene"] == "cube":
            assert np.array_equal(cn, c["color_full"]), (tag, key, "default look colour")
        # in/out default look on pinned buffers: cleared on the host first
        cn[...] = 0; tn[...] = np.inf
        host_call(s, full_frame(c), color, t)
        assert util.same_bits(tn, r["t"]) and np.array_equal(cn, r["color"]), (tag, key, "default look, pinned in/out")
    print(tag, key, "ok")

KEYS = ("example_grid|default|160x90", "sdf_6|default|160x90", "stanford-bunny|default|160x90", "cube|default|128x128",
        "sdf_6|orbit23tilt|96x64", "example_grid|close|96x64")
"""


def _run(code, env=None, timeout=900):
    e = dict(os.environ, **(env or {}))
    r = subprocess.run([sys.executable, "-c", PRELUDE + code], env=e, cwd=ROOT, capture_output=True, text=True,
                       timeout=timeout)
    assert r.returncode == 0 and "all ok" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    return r.stdout


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def test_single_device_host_paths_cover_every_scene_kind_and_look():
    """Kernel-stores-to-host for secondary rays, prim ids, the small-mesh kernel and the in/out contract (one GPU)."""
    _run("""
for key in KEYS:
    check_scene(key, util.cuda_scene(g, g.case(key)["scene"]), "1 gpu")
a = g.scene_arrays("cube")
check_scene("cube|default|128x128", rt.TriangleMesh(a["vpos4"], a["indices"], accel=rt.MESH_BRUTE), "1 gpu brute")
print("all ok")
""")


@pytest.mark.parametrize("band_rows", ["8", "20"])
def test_one_call_renders_on_every_listed_device(band_rows):
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    _run("""
n = torch.cuda.device_count()
rt.init(list(range(n)))
for key in KEYS:
    s = util.cuda_scene(g, g.case(key)["scene"])
    assert s.replicas() == list(range(n)), s.replicas()
    check_scene(key, s, f"{n} gpus")
a = g.scene_arrays("cube")
check_scene("cube|default|128x128", rt.TriangleMesh(a["vpos4"], a["indices"], accel=rt.MESH_BRUTE), f"{n} gpus brute")
# a frame whose height is not a multiple of the band: ragged last band, odd number of bands
c = g.case("example_grid|default|160x90")
s = util.cuda_scene(g, "example_grid")
one = None
for H in (90, 93, 7, 1):
    W = 160
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    r = s.render(f, want_prim=True)
    # the same frame from a single-device call: render the bands one by one on this process' current device
    parts_c = np.zeros((H, W), np.uint32); parts_t = np.full((H, W), np.inf, np.float32)
    for j in range(3):
        q = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, bands=(8, 3, j)), color=parts_c, t=parts_t)
        parts_c, parts_t = q["color"], q["t"]
    assert util.same_bits(r["t"], parts_t) and np.array_equal(r["color"], parts_c), H
    color, t = pinned(H, W, torch.int32, 5), pinned(H, W, torch.float32, 0.5)
    host_call(s, rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True), color, t)
    assert util.same_bits(t.numpy().reshape(H, W), r["t"]) and np.array_equal(color.numpy().view(np.uint32).reshape(H, W), r["color"]), H
# device-resident framebuffer on each device in turn: peers store their bands into it over NVLink
for key in KEYS:
    c = g.case(key)
    s = util.cuda_scene(g, c["scene"])
    W, H = c["W"], c["H"]
    for owner in range(n):
        dev = torch.device("cuda", owner)
        with torch.cuda.device(dev):
            color = torch.full((H * W,), 0x5A5A5A5A, dtype=torch.int32, device=dev)
            t = torch.full((H * W,), 0.25, dtype=torch.float32, device=dev)
            st = torch.cuda.current_stream(dev)
            s.render_device(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True), color.data_ptr(), None,
                            t.data_ptr(), stream=st.cuda_stream)
            st.synchronize()
            assert util.same_bits(t.cpu().numpy().reshape(H, W), c["t"]), (key, owner, "device fan-out depth")
            assert np.array_equal(color.cpu().numpy().view(np.uint32).reshape(H, W), c["color_normal"]), (key, owner)
            # in/out on the device: left half cleared, right half holds a nearer surface
            cv, tv = color.view(H, W), t.view(H, W)
            cv[:, :W // 2] = 0; tv[:, :W // 2] = float("inf"); cv[:, W // 2:] = 7; tv[:, W // 2:] = 1e-3
            s.render_device(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H), color.data_ptr(), t.data_ptr(),
                            t.data_ptr(), stream=st.cuda_stream)
            st.synchronize()
            tn, cn = t.cpu().numpy().reshape(H, W), color.cpu().numpy().view(np.uint32).reshape(H, W)
            assert util.same_bits(tn[:, :W // 2], c["t"][:, :W // 2]) and np.array_equal(cn[:, :W // 2], c["color_normal"][:, :W // 2]), (key, owner, "in/out")
            assert np.all(cn[:, W // 2:] == 7) and np.all(tn[:, W // 2:] == np.float32(1e-3)), (key, owner)
    print("device fan-out", key, "ok")
rt.init([])
s = util.cuda_scene(g, "example_grid")
assert len(s.replicas()) == 1
print("all ok")
""", env={"RT_MULTI_BAND_ROWS": band_rows})


def test_cpp_renderer_draw_on_every_listed_device(tmp_path):
    """hostapi/example_frame_loop (the viewer's frame loop, reference class names) with RT_DEVICES listing every GPU:
    Renderer::draw -> one rt_render_host call -> the reference's frames, pageable and pinned FrameBuffer."""
    import numpy as np
    from tests import util
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (gpurun --gpus 2)")
    exe = os.path.join(ROOT, "triangles_sdf_cpu_raytracing_b200", "hostapi", "example_frame_loop")
    if not os.path.exists(exe):
        pytest.skip("example_frame_loop has not been built (make)")
    g = util.Golden()
    c = g.case("example_grid|default|160x90")
    v = g.scene_arrays("example_grid")["values"]
    with open(tmp_path / "g.grid", "wb") as f:
        f.write(np.array(v.shape, np.uint32).tobytes())
        f.write(np.ascontiguousarray(v, np.float32).tobytes())
    env = dict(os.environ, RT_DEVICES=",".join(str(d) for d in range(n)))
    for extra, look, want_c, want_t in (([], ("normal", "0"), c["color_normal"], c["t"]),
                                        (["0", "pinned"], ("normal", "0"), c["color_normal"], c["t"]),
                                        (["0", "pinned"], ("lambert", "1"), c["color_full"], c["t_full"])):
        out = tmp_path / "o.raw"
        p = subprocess.run([exe, str(tmp_path / "g.grid"), "160", "90", *look, str(out), *extra], env=env,
                           capture_output=True, text=True, timeout=120)
        assert p.returncode == 0, p.stderr
        raw = np.fromfile(out, dtype=np.uint8)
        color = raw[:4 * 160 * 90].view(np.uint32).reshape(90, 160)
        t = raw[4 * 160 * 90:].view(np.float32).reshape(90, 160)
        assert util.same_bits(t, want_t) and np.array_equal(color, want_c), (extra, look)
