"""GPU: rt_intersect_host_static (include/rt_b200.h), the single-ray call for callers that must not allocate.

It has no counterpart in the reference: rt_intersect_host with its device
staging in one fixed file-scope device array instead of buffers allocated per call.  Checked through the C ABI against
rt_intersect_host on the same rays, bit for bit (both run the same kernel in RT_MATH_EXACT), on two calls in a row with
DIFFERENT ray sets of different sizes (the array keeps the last call's contents), on every scene kind, and that a
batch beyond the stated maximum is refused without touching anything.  Collected last on
purpose."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_RAYS = 65536  # RT_INTERSECT_STATIC_MAX_RAYS


def _rays(n, seed, inside=False):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(n, 3))
    o = (v / np.linalg.norm(v, axis=1, keepdims=True) * (0.4 if inside else 2.5)).astype(np.float32)
    target = rng.uniform(-0.6, 0.6, size=(n, 3))
    d = target - o
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return np.ascontiguousarray(o), np.ascontiguousarray(d)


def _call(fn, scene, o, d, plane_y=None):
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    n = o.shape[0]
    hit = np.full(n, 0x77, np.uint8)
    t = np.full(n, -5.0, np.float32)
    nrm = np.full((n, 3), -5.0, np.float32)
    prim = np.full(n, 0x12345678, np.uint32)
    p = lambda a: C.c_void_p(a.ctypes.data)
    rc = fn(scene.handle, p(o), p(d), n, 0.01, 100.0, int(plane_y is not None), float(plane_y or 0.0), p(hit), p(t),
            p(nrm), p(prim))
    return rc, hit, t, nrm, prim


def _same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


@pytest.mark.parametrize("name", ("example_grid", "sdf_6", "stanford-bunny", "cube"))
def test_static_scratch_call_equals_the_allocating_call_twice_in_a_row(name):
    import triangles_sdf_cpu_raytracing_b200 as rt
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    from tests import util
    rt.set_math_mode(rt.MATH_EXACT)
    g = util.Golden()
    s = util.cuda_scene(g, name)
    lib = L.lib()
    hits = 0
    # second call: fewer rays than the first, other rays, from inside the model's box, with the ground plane
    for n, seed, inside, plane_y in ((5000, 1, False, None), (1237, 2, True, -1.0)):
        o, d = _rays(n, seed, inside)
        rc0, *want = _call(lib.rt_intersect_host, s, o, d, plane_y)
        rc1, *got = _call(lib.rt_intersect_host_static, s, o, d, plane_y)
        assert rc0 == 0 and rc1 == 0, (rc0, rc1, lib.rt_last_error().decode())
        for w, x, what in zip(want, got, ("hit", "t", "normal", "prim")):
            assert _same(w, x), f"{name}: {what} differs on {(w != x).sum()} of {n} rays (call with seed {seed})"
        hits += int(want[0].sum())
        assert set(np.unique(want[0])) <= {0, 1}
    assert hits > 0
    s.destroy()


def test_static_scratch_call_at_its_maximum_and_beyond():
    import triangles_sdf_cpu_raytracing_b200 as rt
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    from tests import util
    rt.set_math_mode(rt.MATH_EXACT)
    g = util.Golden()
    s = util.cuda_scene(g, "example_grid")
    lib = L.lib()
    o, d = _rays(MAX_RAYS + 1, 3)
    rc, hit, t, nrm, prim = _call(lib.rt_intersect_host_static, s, o, d)
    assert rc != 0, "a batch beyond RT_INTERSECT_STATIC_MAX_RAYS must be refused"
    msg = lib.rt_last_error().decode()
    assert str(MAX_RAYS) in msg and "static" in msg, msg
    assert np.all(hit == 0x77) and np.all(t == np.float32(-5.0)) and np.all(prim == 0x12345678)  # nothing was written
    # exactly the maximum works, and still equals the allocating call
    o, d = o[:MAX_RAYS], d[:MAX_RAYS]
    rc0, *want = _call(lib.rt_intersect_host, s, np.ascontiguousarray(o), np.ascontiguousarray(d))
    rc1, *got = _call(lib.rt_intersect_host_static, s, np.ascontiguousarray(o), np.ascontiguousarray(d))
    assert rc0 == 0 and rc1 == 0
    assert all(_same(w, x) for w, x in zip(want, got))
    # n = 0 is a no-op, a null scene an error (no crash)
    assert lib.rt_intersect_host_static(s.handle, None, None, 0, 0.01, 100.0, 0, 0.0, None, None, None, None) == 0
    assert lib.rt_intersect_host_static(None, None, None, 1, 0.01, 100.0, 0, 0.0, None, None, None, None) != 0
    s.destroy()
