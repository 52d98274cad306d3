#!/usr/bin/env python
"""Kernel time only (device-timed, L2 flushed between launches) of bench workloads, for A/B runs of kernel variants:

    [RT_B200_LIB=path] python tools/ktime.py [--sched S] [--reps N] [--check] scene ...

--check compares the frame with the plain-C oracle at 480x270 (bit-exact or abort)."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import triangles_sdf_cpu_raytracing_b200 as rt  # noqa: E402
from tests import util  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("scenes", nargs="+")
ap.add_argument("--sched", default="auto")
ap.add_argument("--reps", type=int, default=30)
ap.add_argument("--math", default="exact")
ap.add_argument("--check", action="store_true")
ap.add_argument("--bands", type=int, default=1, help="time rank 0's share of a frame split into 8-row bands over this many GPUs")
args = ap.parse_args()
rt.set_math_mode(rt.MATH_FAST if args.math == "fast" else rt.MATH_EXACT)
rt.set_sched_mode({"auto": rt.SCHED_AUTO, "tiles": rt.SCHED_TILES, "persistent": rt.SCHED_PERSISTENT, "compact": rt.SCHED_COMPACT}[args.sched])
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
tag = os.path.basename(os.environ.get("RT_B200_LIB", "default"))
for name in args.scenes:
    g, scene_name, c, W, H, look = bench.load_workload(name)
    arrays = g.scene_arrays(scene_name)
    lk, plane_y = bench.look_kwargs(look, bench.model_plane_y(arrays))
    mode = rt.SHADE_LAMBERT if look == "default" else rt.SHADE_NORMAL
    s = util.cuda_scene(g, scene_name)
    f = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=mode, plane_y=plane_y, clear=True,
                      bands=(8, args.bands, 0) if args.bands > 1 else None, **lk)
    color = torch.zeros(W * H, dtype=torch.int32, device=dev)
    t = torch.zeros(W * H, dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream()
    ok = ""
    if args.check:
        w, h = 480, 270
        if abs(w / h - W / H) > 1e-6:
            w, h = 256, 256
        r = s.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], w, h, mode=mode, plane_y=plane_y, **lk))
        want = bench.oracle_frame(g, scene_name, c, w, h, look)
        same = want is not None and util.same_bits(r["t"], want[1]) and np.array_equal(r["color"], want[0])
        ok = " bit-exact" if same else " MISMATCH"
    for _ in range(3):
        s.render_device(f, color.data_ptr(), None, t.data_ptr(), stream=st.cuda_stream)
    torch.cuda.synchronize()
    ev = []
    for _ in range(args.reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s.render_device(f, color.data_ptr(), None, t.data_ptr(), stream=st.cuda_stream)
        b.record()
        ev.append((a, b))
    torch.cuda.synchronize()
    ms = [a.elapsed_time(b) for a, b in ev]
    print(f"{tag:28s} {name:28s} 1/{args.bands} sched {args.sched:10s} mean {np.mean(ms):8.4f} ms  min {np.min(ms):8.4f}  {W * H / args.bands / np.mean(ms) / 1e3:9.0f} Mrays/s{ok}",
          flush=True)
    s.destroy()
