#!/usr/bin/env python
"""One process, n GPUs (rt_init): the same single rt_render_host / rt_render_device call at n = 1, 2, 4, 8.

    python tools/multi_e2e.py [scene ...]          (RT_MULTI_BAND_ROWS=r in the environment selects the band height)

Per scene and n: wall-clock ms of the host call on pinned buffers (fused clear / in-out) and on pageable buffers, the
slowest GPU's kernel ms, and the device-timed ms of the device-resident call (bands gathered into GPU 0's HBM over NVLink
inside the render kernels).  Every frame is compared with the n = 1 frame (bit-identical or the run aborts)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import triangles_sdf_cpu_raytracing_b200 as rt  # noqa: E402
from tests import util  # noqa: E402
from triangles_sdf_cpu_raytracing_b200 import _lib as L  # noqa: E402

def numa_interleave():
    """RT_NUMA=interleave: page-locked buffers allocated from here on are spread page by page over every NUMA node
    (set_mempolicy(MPOL_INTERLEAVE)); an experiment about what limits n GPUs storing into ONE host framebuffer."""
    import glob
    nodes = len(glob.glob("/sys/devices/system/node/node[0-9]*"))
    if nodes < 2:
        return f"{nodes} NUMA node(s): nothing to interleave"
    libc = C.CDLL(None, use_errno=True)
    mask = C.c_ulong((1 << nodes) - 1)
    rc = libc.syscall(238, 3, C.byref(mask), C.c_ulong(65))  # SYS_set_mempolicy, MPOL_INTERLEAVE
    return f"{nodes} NUMA nodes, set_mempolicy(MPOL_INTERLEAVE) rc={rc} errno={C.get_errno()}"


if os.environ.get("RT_NUMA") == "interleave":
    print("#", numa_interleave(), flush=True)
scenes = sys.argv[1:] or ["grid513_4k", "grid65_1080p", "octree8_4k", "bunny_4k", "default_look_1080p"]
avail = torch.cuda.device_count()
reps = int(os.environ.get("REPS", "20"))
rows = []
for name in scenes:
    g, scene_name, c, W, H, look = bench.load_workload(name)
    arrays = g.scene_arrays(scene_name)
    lk, plane_y = bench.look_kwargs(look, bench.model_plane_y(arrays))
    mode = rt.SHADE_LAMBERT if look == "default" else rt.SHADE_NORMAL
    px = W * H
    ref = None
    for n in (1, 2, 4, 8):
        if n > avail:
            break
        rt.init(list(range(n)))
        s = util.cuda_scene(g, scene_name)
        mk = lambda **kw: rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, mode=mode, plane_y=plane_y, **lk, **kw)  # noqa: E731
        hc = torch.full((px,), 0x5A5A5A5A, dtype=torch.int32).pin_memory()
        ht = torch.full((px,), 0.25, dtype=torch.float32).pin_memory()
        dms, tms = C.c_float(), C.c_float()

        def call(fr, cb, tb):
            L.check(L.lib().rt_render_host(s.handle, C.byref(fr), C.c_void_p(cb.data_ptr()), C.c_void_p(tb.data_ptr()), None,
                                           C.byref(dms), C.byref(tms), None))

        def timed(fr, cb, tb, before=lambda: None, k=reps):
            ts, ks = [], []
            for i in range(k + 3):
                before()
                t0 = time.perf_counter()
                call(fr, cb, tb)
                if i >= 3:
                    ts.append((time.perf_counter() - t0) * 1e3)
                    ks.append(dms.value)
            return float(np.mean(ts)), float(np.min(ts)), float(np.mean(ks))
        clear = mk(clear=True)
        call(clear, hc, ht)
        frame = (hc.numpy().view(np.uint32).reshape(H, W).copy(), ht.numpy().reshape(H, W).copy())
        if ref is None:
            ref = frame
        assert np.array_equal(frame[0], ref[0]) and util.same_bits(frame[1], ref[1]), (name, n, "host frame differs from n = 1")
        e_ms, e_min, k_ms = timed(clear, hc, ht)

        def hclear():
            hc.zero_()
            ht.fill_(float("inf"))
        io_ms, _, _ = timed(mk(), hc, ht, hclear)
        assert np.array_equal(hc.numpy().view(np.uint32).reshape(H, W), ref[0]), (name, n, "in/out frame differs")
        pc, pt = torch.zeros(px, dtype=torch.int32), torch.full((px,), float("inf"), dtype=torch.float32)

        def pclear():
            pc.zero_()
            pt.fill_(float("inf"))
        pg_ms, _, _ = timed(mk(), pc, pt, pclear, k=max(3, reps // 4))
        assert np.array_equal(pc.numpy().view(np.uint32).reshape(H, W), ref[0]), (name, n, "pageable frame differs")
        # device-resident framebuffer on GPU 0, asynchronous call, CUDA events around it
        dev = torch.device("cuda", 0)
        with torch.cuda.device(dev):
            dc = torch.zeros(px, dtype=torch.int32, device=dev)
            dt = torch.zeros(px, dtype=torch.float32, device=dev)
            flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
            st = torch.cuda.current_stream(dev)
            for _ in range(3):
                s.render_device(clear, dc.data_ptr(), None, dt.data_ptr(), stream=st.cuda_stream)
            st.synchronize()
            ev = []
            for _ in range(reps):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                s.render_device(clear, dc.data_ptr(), None, dt.data_ptr(), stream=st.cuda_stream)
                b.record()
                ev.append((a, b))
            st.synchronize()
            d_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
            assert np.array_equal(dc.cpu().numpy().view(np.uint32).reshape(H, W), ref[0]), (name, n, "device frame differs")
            assert util.same_bits(dt.cpu().numpy().reshape(H, W), ref[1]), (name, n, "device depth differs")
        row = dict(scene=name, n=n, band_rows=int(os.environ.get("RT_MULTI_BAND_ROWS", "8")), e2e_ms=e_ms, e2e_min_ms=e_min,
                   slowest_kernel_ms=k_ms, inout_ms=io_ms, pageable_ms=pg_ms, device_gather_ms=d_ms,
                   e2e_mrays=px / e_ms / 1e3, device_mrays=px / d_ms / 1e3)
        rows.append(row)
        print(json.dumps(row), flush=True)
        s.destroy()
rt.init([])
print(f"\n{'scene':22s} {'n':>2s} {'band':>4s} {'e2e ms':>8s} {'min':>8s} {'kern':>8s} {'in/out':>8s} {'pageable':>8s} {'dev ms':>8s} {'e2e Mray/s':>10s} {'dev Mray/s':>10s}")
for r in rows:
    print(f"{r['scene']:22s} {r['n']:2d} {r['band_rows']:4d} {r['e2e_ms']:8.3f} {r['e2e_min_ms']:8.3f} {r['slowest_kernel_ms']:8.3f} "
          f"{r['inout_ms']:8.3f} {r['pageable_ms']:8.3f} {r['device_gather_ms']:8.3f} {r['e2e_mrays']:10.0f} {r['device_mrays']:10.0f}")
