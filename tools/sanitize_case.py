"""Small invocation of every kernel family for compute-sanitizer (memcheck): all scene kinds, every scheduling mode,
secondary rays, counters, the device BVH builder, on-device ingestion, brick layout, single rays.
   compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import triangles_sdf_cpu_raytracing_b200 as rt
from tests import util

g = util.Golden()
n = 0
for key in ("example_grid|default|160x90", "sdf_6|default|160x90", "stanford-bunny|default|160x90", "cube|default|128x128"):
    c = g.case(key)
    scenes = [util.cuda_scene(g, c["scene"])]
    if c["scene"] == "example_grid":
        scenes.append(util.cuda_scene(g, c["scene"], layout=rt.GRID_BRICKS))
    if c["scene"] == "stanford-bunny":
        scenes.append(util.cuda_scene(g, c["scene"], accel=rt.MESH_LBVH_GPU))
    for s in scenes:
        for sched in (rt.SCHED_TILES, rt.SCHED_PERSISTENT, rt.SCHED_COMPACT):
            rt.set_sched_mode(sched)
            for clear in (False, True):
                f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], clear=clear)
                r = s.render(f, want_prim=True, want_counters=True)
                assert util.same_bits(r["t"], c["t"]), key
                r = s.render(f)
                assert util.same_bits(r["t"], c["t"]), key
                n += 2
        rt.set_sched_mode(rt.SCHED_AUTO)
        f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], c["W"], c["H"], mode=rt.SHADE_LAMBERT, shadows=True,
                          reflections=True, plane_y=-1.0)
        s.render(f)
        f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], 37, 23, bands=(7, 3, 1))
        s.render(f)
        o = np.tile(np.asarray(c["rayPos"], np.float32), (64, 1))
        d = np.random.default_rng(0).normal(size=(64, 3)).astype(np.float32)
        s.intersect(o, d / np.linalg.norm(d, axis=1, keepdims=True))
        n += 3
        s.destroy()
print(f"sanitize_case ok: {n} launches checked")
