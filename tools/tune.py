#!/usr/bin/env python
"""Measurement sweeps over the library's tuning knobs (RT_TILE_BLOCK, RT_HOST_CHUNKS are read once at load, so
every point is its own process).   python tools/tune.py sweep   |   python tools/tune.py point SCENE SCHED"""
import ctypes as C, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def point(scene_name, sched):
    import numpy as np, torch
    import triangles_sdf_cpu_raytracing_b200 as rt
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    import bench
    g, sname, c, W, H = bench.load_workload(scene_name)
    rt.set_sched_mode({"auto": rt.SCHED_AUTO, "tiles": rt.SCHED_TILES, "persistent": rt.SCHED_PERSISTENT, "compact": rt.SCHED_COMPACT}[sched])
    from tests import util
    scene = util.cuda_scene(g, sname)
    frame = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, clear=True)
    px = W * H
    color = torch.zeros(px, dtype=torch.int32, device="cuda")
    tbuf = torch.zeros(px, dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(5):
        scene.render_device(frame, color.data_ptr(), None, tbuf.data_ptr(), stream=st)
    torch.cuda.synchronize()
    ev = []
    for _ in range(30):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); scene.render_device(frame, color.data_ptr(), None, tbuf.data_ptr(), stream=st); b.record()
        ev.append((a, b))
    torch.cuda.synchronize()
    k = [a.elapsed_time(b) for a, b in ev]
    out = {"scene": scene_name, "sched": sched, "tile_block": os.environ.get("RT_TILE_BLOCK", "-"),
           "chunks": os.environ.get("RT_HOST_CHUNKS", "-"), "kernel_ms": round(float(np.mean(k)), 4),
           "kernel_ms_min": round(float(np.min(k)), 4)}
    if os.environ.get("TUNE_E2E"):
        hc = torch.zeros(px, dtype=torch.int32).pin_memory()
        ht = torch.zeros(px, dtype=torch.float32).pin_memory()
        dms, tms = C.c_float(), C.c_float()
        sink = [0]
        modes = {"untouched": lambda: None, "cpu_read": lambda: sink.__setitem__(0, sink[0] + int(hc[::16].sum())),
                 "cpu_write": lambda: (hc.fill_(0x5A5A5A5A), ht.fill_(0.25))}
        for name, before in modes.items():
            ts = []
            for i in range(23):
                before()
                t0 = time.perf_counter()
                L.check(L.lib().rt_render_host(scene.handle, C.byref(frame), C.c_void_p(hc.data_ptr()),
                                               C.c_void_p(ht.data_ptr()), None, C.byref(dms), C.byref(tms), None))
                if i >= 3:
                    ts.append((time.perf_counter() - t0) * 1e3)
            out["e2e_ms_" + name] = round(float(np.mean(ts)), 4)
        out["e2e_total_ms_event"] = round(tms.value, 4)
        # zero copy: the kernel stores straight into the pinned host framebuffer (UVA: host pointer == device pointer)
        ts = []
        for i in range(23):
            t0 = time.perf_counter()
            scene.render_device(frame, hc.data_ptr(), None, ht.data_ptr(), stream=st)
            torch.cuda.synchronize()
            if i >= 3:
                ts.append((time.perf_counter() - t0) * 1e3)
        out["zerocopy_ms"] = round(float(np.mean(ts)), 4)
    print(json.dumps(out), flush=True)


def run(env, *args):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, __file__, "point", *args], env=e, capture_output=True, text=True)
    sys.stdout.write(r.stdout if r.returncode == 0 else f"FAILED {env} {args}: {r.stderr[-400:]}\n")
    sys.stdout.flush()


if __name__ == "__main__":
    if sys.argv[1] == "point":
        point(sys.argv[2], sys.argv[3])
    else:
        for ch in ("0", "1", "4"):
            run({"RT_HOST_CHUNKS": ch, "TUNE_E2E": "1"}, "grid65_1080p", "auto")
        for tb in ("32", "64", "128", "256"):
            run({"RT_TILE_BLOCK": tb, "TUNE_E2E": "1"}, "grid65_1080p", "tiles")
        run({"TUNE_E2E": "1"}, "grid65_4k", "auto")
        run({"TUNE_E2E": "1"}, "octree6_4k", "auto")
        run({"TUNE_E2E": "1"}, "cube_512", "auto")
