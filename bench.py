#!/usr/bin/env python
"""bench.py — Mrays/s of the primary-ray hot path on B200, next to the reference's CPU renderer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--scene NAME] [--math exact|fast]

A "step" is one frame: every primary ray of the workload goes through the CUDA path once (clear + render, and for
N > 1 the gather of the screen-space tiles to rank 0).  The N=1 workload is BASELINE.json configs[1]: the
stanford-bunny dense SDF grid (example_grid.grid, 65^3) sphere-traced at 1920x1080 with the viewer's start
camera, ShadingMode::Normal, model only.  Prints ONE JSON line (see DESIGN.md "Measurement").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tests import util  # noqa: E402  (fixtures only: scene arrays + the camera matrices the reference computed)

WORKLOADS = {
    # name: (scene fixture, golden case holding the camera matrices for this aspect ratio, W, H)
    "grid65_1080p": ("example_grid", "example_grid|default|160x90", 1920, 1080),
    "cube_512": ("cube", "cube|default|128x128", 512, 512),
    "octree6_4k": ("sdf_6", "sdf_6|default|160x90", 3840, 2160),
    "bunny_4k": ("stanford-bunny", "stanford-bunny|default|160x90", 3840, 2160),
    "grid65_4k": ("example_grid", "example_grid|default|160x90", 3840, 2160),
    # deterministic stand-ins for the resources the reference tree does not ship (standins.py)
    "grid257_4k": ("example_grid_x4", "example_grid|default|160x90", 3840, 2160),   # configs[2], 68 MB: L2-resident
    "grid513_4k": ("example_grid_x8", "example_grid|default|160x90", 3840, 2160),   # configs[2], 540 MB: HBM-resident
    "octree8_4k": ("octree_d8", "example_grid|default|160x90", 3840, 2160),         # configs[3]
    "octree9_4k": ("octree_d9", "example_grid|default|160x90", 3840, 2160),         # configs[3]
    "bunny_s2_4k": ("stanford-bunny_s2", "example_grid|default|160x90", 3840, 2160),  # configs[4] mesh, 1.1 M triangles
    # configs[4]: step k renders frame k % 64 of the orbit (Camera::rotate(2*pi/64/0.01, 0) k times, camera.cpp:5-20)
    "bunny_s2_orbit64_4k": ("stanford-bunny_s2", "example_grid|default|160x90", 3840, 2160),
}
ORBIT_DX = float(np.float32(2.0 * np.pi / 64.0 / 0.01))
HEADLINE = "grid65_1080p"
METRIC = "Mrays/s primary rays"
L2_FLUSH_BYTES = 256 << 20


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.06)
            self.proc.terminate()
            self.th.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def load_workload(name):
    g = util.Golden()
    scene_name, case_key, W, H = WORKLOADS[name]
    c = g.case(case_key)
    assert abs(W / H - c["W"] / c["H"]) < 1e-6
    return g, scene_name, c, W, H


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref when it was compiled, else the C port),
    all host threads, on the same workload.  Each step = a bounded band of rows of the frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    g, scene_name, c, W, H = load_workload(args.scene)
    from oracle import ref_oracle as R
    cores = os.cpu_count() or 1
    if R.available(contract=True):
        o = R.RefOracle(contract=True)  # GCC's default contraction = the reference's Release build
        o.set_threads(cores)
        a = g.scene_arrays(scene_name)
        s = o.grid(a["values"]) if "values" in a else o.octree(a["nodes36"]) if "nodes36" in a else o.mesh(a["vpos4"], a["indices"])
        kind = "reference"
        orbit = [0]

        def rows(y0, y1):
            cam = R.CamSpec(n_rotate=orbit[0] % 64, dx=ORBIT_DX, dy=0.0) if "orbit" in args.scene else R.CamSpec()
            orbit[0] += 1
            return s.render_rows(cam, W, H, y0, y1, proj_inv=c["projInv"])[2]
    else:
        from oracle import port_oracle as P
        s = util.port_scene(g, scene_name)
        kind = "port"

        def rows(y0, y1):
            t0 = time.perf_counter()
            s.render(c["rayPos"], c["viewInv"], c["projInv"], W, H, rows=(y0, y1))
            return (time.perf_counter() - t0) * 1e3
    # size the per-step sample: a centred band of rows (the model is mid-screen) so that the run stays bounded
    probe_rows = max(8, H // 16)
    y0 = (H - probe_rows) // 2
    ms_probe = rows(y0, y0 + probe_rows)
    budget_ms = 90_000.0 / max(1, args.steps + args.warmup)
    n_rows = int(min(H, max(8, probe_rows * budget_ms / max(ms_probe, 1e-3))))
    y0 = (H - n_rows) // 2
    for _ in range(args.warmup):
        rows(y0, y0 + n_rows)
    total = 0.0
    for _ in range(args.steps):
        total += rows(y0, y0 + n_rows)
    ms = total / args.steps
    value = n_rows * W / ms / 1e3
    sample = f"rows {y0}..{y0 + n_rows} of {H} ({n_rows * W} rays per step), Renderer-equivalent row loop, OpenMP dynamic"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "reference scene file (fixture copy), viewer start camera",
        "config": {"workload": args.scene, "scene": scene_name, "width": W, "height": H, "shading": "Normal"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ CUDA arm
def cpu_baseline(g, scene_name, c, W, H, orbit=False):
    """Rank 0, N=1: the reference's CPU path timed on this box's host cores on a bounded sample."""
    from oracle import ref_oracle as R
    cores = os.cpu_count() or 1
    try:
        if R.available(contract=True):
            o = R.RefOracle(contract=True)
            o.set_threads(cores)
            a = g.scene_arrays(scene_name)
            s = o.grid(a["values"]) if "values" in a else o.octree(a["nodes36"]) if "nodes36" in a else o.mesh(a["vpos4"], a["indices"])
            if orbit:
                ks = (0, 16, 32, 48)
                ms = [s.render_timed(R.CamSpec(n_rotate=k, dx=ORBIT_DX, dy=0.0), W, H, frames=2, proj_inv=c["projInv"])[0]
                      for k in ks]
                return {"value": W * H / float(np.mean(ms)) / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                        "sample": f"orbit frames {ks} of 64 at {W}x{H} through the reference's Renderer::draw (best of 2 "
                                  f"each, mean over the 4 cameras), scalar ISPC stand-in"}
            best, each = s.render_timed(R.CamSpec(), W, H, frames=3, proj_inv=c["projInv"])
            return {"value": W * H / best / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                    "sample": f"3 full {W}x{H} frames through the reference's Renderer::draw (best), scalar ISPC stand-in"}
    except Exception as e:  # noqa: BLE001 - fall through to the port
        sys.stderr.write(f"reference oracle unavailable ({e}); timing the C port\n")
    s = util.port_scene(g, scene_name)
    t0 = time.perf_counter()
    s.render(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    ms = (time.perf_counter() - t0) * 1e3
    return {"value": W * H / ms / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": f"1 full {W}x{H} frame through oracle/port_oracle.c (OpenMP)"}


def oracle_frame(g, scene_name, c, W, H):
    """The checker's frame for the timed workload (outside any timing). SDF scenes: the plain-C oracle. Meshes: the
    reference's own code when oracle/_ref is present (it builds its BVH); the C oracle's brute force would need
    rays x triangles tests and is only used for small meshes."""
    a = g.scene_arrays(scene_name)
    if "vpos4" not in a or a["indices"].size // 3 <= 64:
        wc, wt, _, _ = util.port_scene(g, scene_name).render(c["rayPos"], c["viewInv"], c["projInv"], W, H)
        return wc, wt, "oracle/port_oracle.c"
    from oracle import ref_oracle as R
    if not R.available():
        return None
    o = R.RefOracle()
    wc, wt, _ = o.mesh(a["vpos4"], a["indices"]).render(R.CamSpec(), W, H, proj_inv=c["projInv"])
    return wc, wt, "oracle/_ref (reference C++, -ffp-contract=off)"


def algorithmic_work(scene_kind, cnt, rays):
    """algorithmic bytes and flops per frame from the live workload counters."""
    S = cnt["grid_samples"] / rays
    I, Lf, Ls = cnt["oct_inner"] / rays, cnt["oct_leaf"] / rays, cnt["oct_leaf_samples"] / rays
    BI, BL, BT = cnt["bvh_inner"] / rays, cnt["bvh_leaf"] / rays, cnt["bvh_tris"] / rays
    hit = cnt["hits"] / rays
    if scene_kind == "grid":
        b, f = 32 * S + 8, 150 + 60 * S
    elif scene_kind == "octree":
        b, f = 36 * (I + Lf) + 8, 107 + 300 * I + 70 * Lf + 73 * Ls + 130 * hit
    else:
        b, f = 204 * BI + 8 * BL + 60 * BT + 8, 107 + 254 * BI + 90 * BT
    return b * rays, f * rays


def run_cuda(args):
    import torch
    import torch.distributed as dist

    import triangles_sdf_cpu_raytracing_b200 as rt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rt.set_math_mode(rt.MATH_FAST if args.math == "fast" else rt.MATH_EXACT)
    rt.set_sched_mode({"auto": rt.SCHED_AUTO, "tiles": rt.SCHED_TILES, "persistent": rt.SCHED_PERSISTENT, "compact": rt.SCHED_COMPACT}[args.sched])

    g, scene_name, c, W, H = load_workload(args.scene)
    t_build = time.perf_counter()
    layout_kw = {}
    arrays = g.scene_arrays(scene_name)  # (stand-ins are generated here, outside the build stopwatch)
    t_build = time.perf_counter()
    if "values" in arrays and args.grid_layout != "auto":
        layout_kw = {"layout": {"linear": rt.GRID_LINEAR, "bricks": rt.GRID_BRICKS}[args.grid_layout]}
    if "vpos4" in arrays and args.mesh_accel != "auto":
        layout_kw = {"accel": {"host": rt.MESH_BVH8, "lbvh_gpu": rt.MESH_LBVH_GPU}[args.mesh_accel]}
    scene = util.cuda_scene(g, scene_name, **layout_kw)
    t_build = time.perf_counter() - t_build
    kind = {0: "grid", 1: "octree", 2: "mesh"}[scene.kind]
    band_rows = args.band_rows
    bands = (band_rows, world, rank) if world > 1 else None
    # FrameBuffer::clear (raytracing.hpp:16-19) is fused into the render launch: one kernel writes every pixel once
    frame = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, bands=bands, clear=True)
    frames = [frame]
    orbit = "orbit" in args.scene
    if orbit:  # the 64 cameras of the orbit, from the host mirror of the reference's Camera (camera.cpp:5-20)
        cams = [rt.host_camera(n_rotate=k, dx=ORBIT_DX, dy=0.0) for k in range(64)]
        frames = [rt.make_frame(pos, vinv, c["projInv"], W, H, bands=bands, clear=True) for pos, _v, vinv in cams]
    step_no = [0]
    px = W * H
    color = torch.zeros(px, dtype=torch.int32, device=dev)
    tbuf = torch.full((px,), float("inf"), dtype=torch.float32, device=dev)
    color_ptr, t_ptr = color.data_ptr(), tbuf.data_ptr()
    shared = None
    if world > 1 and args.gather == "p2p":
        # the root's framebuffer mapped into every rank (CUDA IPC): kernels store their tiles into it over NVLink
        from triangles_sdf_cpu_raytracing_b200.multigpu import SharedFramebuffer
        shared = SharedFramebuffer(W, H, root=0)
        color_ptr, t_ptr = shared.color.value, shared.t.value
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    n_bands = (H + band_rows - 1) // band_rows
    pad_rows = n_bands * band_rows

    def gather_tiles():
        """N > 1: every rank ships its interleaved row bands to rank 0 (NCCL send/recv over NVLink)."""
        if world == 1 or shared is not None:
            return
        cv = color.view(H, W)
        tv = tbuf.view(H, W)
        if rank == 0:
            ops, stage = [], []
            for r in range(1, world):
                idx = [b for b in range(n_bands) if b % world == r]
                nrows = sum(min(band_rows, H - b * band_rows) for b in idx)
                sc = torch.empty((nrows, W), dtype=torch.int32, device=dev)
                st = torch.empty((nrows, W), dtype=torch.float32, device=dev)
                stage.append((idx, sc, st))
                ops += [dist.P2POp(dist.irecv, sc, r), dist.P2POp(dist.irecv, st, r)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()
            for idx, sc, st in stage:
                o = 0
                for b in idx:
                    n = min(band_rows, H - b * band_rows)
                    cv[b * band_rows:b * band_rows + n] = sc[o:o + n]
                    tv[b * band_rows:b * band_rows + n] = st[o:o + n]
                    o += n
        else:
            idx = [b for b in range(n_bands) if b % world == rank]
            sc = torch.cat([cv[b * band_rows:min(H, (b + 1) * band_rows)] for b in idx])
            st = torch.cat([tv[b * band_rows:min(H, (b + 1) * band_rows)] for b in idx])
            for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, sc, 0), dist.P2POp(dist.isend, st, 0)]):
                w.wait()

    def step():
        # FrameBuffer::clear + Renderer::draw = ONE launch of our kernel
        scene.render_device(frames[step_no[0] % len(frames)], color_ptr, None, t_ptr, stream=stream.cuda_stream)
        step_no[0] += 1
        gather_tiles()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    step_no[0] = 0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    clocks = ClockSampler(local)
    clocks.__enter__()  # sampled across all three device-timed loops below
    barrier()
    for a, b in ev:
        flush.zero_()  # evict the scene and the framebuffer from L2 between timed iterations (untimed)
        if world > 1:
            dist.barrier()
        a.record()
        step()
        b.record()
    barrier()
    ms_total = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        tt = torch.tensor([ms_total], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = px / ms_step / 1e3

    # warm-L2 variant (frames back to back, as a viewer would issue them), kernel only
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(args.steps):
        scene.render_device(frames[i % len(frames)], color.data_ptr(), None, tbuf.data_ptr(), stream=stream.cuda_stream)
    b.record()
    barrier()
    kernel_ms_warm = a.elapsed_time(b) / args.steps
    # kernel alone, cold L2 (the roofline's launch duration)
    kev = []
    for i in range(min(args.steps, 64)):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        scene.render_device(frames[i % len(frames)], color.data_ptr(), None, tbuf.data_ptr(), stream=stream.cuda_stream)
        b.record()
        kev.append((a, b))
    barrier()
    clocks.__exit__()
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))

    multi_frame = None
    if world > 1:
        # the gathered frame of the last step, checked on rank 0 against the oracle below
        # (orbit: render frame 0 once more, the camera the checker uses)
        barrier()
        if orbit:
            step_no[0] = 0
            step()
            barrier()
        if rank == 0:
            if shared is not None:
                multi_frame = shared.download()
            else:
                multi_frame = (color.cpu().numpy().view(np.uint32).reshape(H, W), tbuf.cpu().numpy().reshape(H, W))
        barrier()
        if shared is not None:
            shared.close()
    staged = int(rt.lib().rt_staged_launch_count())
    if world > 1:  # ranks != 0 are the ones that store into a peer framebuffer
        tt = torch.tensor([staged], device=dev, dtype=torch.int64)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        staged = int(tt.item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- end to end through the host-buffer C-ABI call (pinned host framebuffer, copies inside the timed region)
    e2e = None
    if world == 1:
        hc = torch.zeros(px, dtype=torch.int32).pin_memory()
        ht = torch.full((px,), float("inf"), dtype=torch.float32).pin_memory()
        import ctypes as C
        from triangles_sdf_cpu_raytracing_b200 import _lib as L
        dms, tms = C.c_float(), C.c_float()
        n_e2e = max(5, min(args.steps, 30))

        def time_host(fr, before):
            """before(i): what the CPU does to the host framebuffer between two calls (untimed, like the reference's own
            stopwatch which only brackets draw).  It matters: cache lines the CPU has just written are dirty and the
            DMA write that follows has to evict them first."""
            times = []
            for i in range(n_e2e + 3):
                before()
                t0 = time.perf_counter()
                L.check(L.lib().rt_render_host(scene.handle, C.byref(fr), C.c_void_p(hc.data_ptr()),
                                               C.c_void_p(ht.data_ptr()), None, C.byref(dms), C.byref(tms), None))
                if i >= 3:
                    times.append((time.perf_counter() - t0) * 1e3)
            return float(np.mean(times))

        def host_clear():   # FrameBuffer::clear() on the host (raytracing.hpp:16-19), main.cpp:197
            hc.zero_()
            ht.fill_(float("inf"))

        def host_dirty():
            hc.fill_(0x5A5A5A5A)
            ht.fill_(0.25)

        sink = [0]

        def host_blit():    # the viewer reads the colour buffer (SDL_UpdateTexture, main.cpp:238-249)
            sink[0] += int(hc[::16].sum())

        # correctness first: dirty buffers, the fused clear must overwrite every pixel (checked against the oracle below)
        host_dirty()
        L.check(L.lib().rt_render_host(scene.handle, C.byref(frame), C.c_void_p(hc.data_ptr()),
                                       C.c_void_p(ht.data_ptr()), None, C.byref(dms), C.byref(tms), None))
        chk_c = hc.numpy().view(np.uint32).reshape(H, W).copy()
        chk_t = ht.numpy().reshape(H, W).copy()
        # (a) clear + draw as one call (rt_frame.clear): only the frame parameters go up, colour + depth come back;
        #     the caller does not touch the buffers between calls
        e2e_ms = time_host(frame, lambda: None)
        blit_ms = time_host(frame, host_blit)
        dirty_ms = time_host(frame, host_dirty)
        # (b) the reference's literal in/out contract: the caller clears on the host, depth + colour go up, both come back
        inout = rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H)
        inout_ms = time_host(inout, host_clear)
        e2e = {"value": px / e2e_ms / 1e3, "unit": "Mrays/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": C.sizeof(L.RtFrame), "d2h_bytes_per_step": 8 * px,
               "api": "rt_render_host, rt_frame.clear=1 (FrameBuffer::clear + Renderer::draw in one call), pinned host "
                      "framebuffer left untouched by the CPU between calls",
               "transfer": ("pipelined copy-engine downloads (RT_HOST_ZEROCOPY=0)"
                            if os.environ.get("RT_HOST_ZEROCOPY", "2") == "0" or W % 8 or args.scene == "cube_512"
                            or args.sched in ("persistent", "compact") else
                            "the render kernel stores its tiles into the pinned host framebuffer over PCIe "
                            "(128-byte bulk stores, one launch, no copy engine)"),
               "d2h_floor_ms_at_56GBps": 8 * px / 56e9 * 1e3,
               # the host call is bound by the PCIe link, not by HBM: achieved = bytes that crossed it / wall time of the
               # call; peak = the copy-engine D2H rate measured on this pool's boxes (tools/pcie_probe.py, Gen5 x16)
               "link": {"bound": "pcie", "achieved": 8 * px / (e2e_ms * 1e-3) / 1e9, "peak": 56.0, "unit": "GB/s",
                        "frac": 8 * px / (e2e_ms * 1e-3) / 1e9 / 56.0},
               "after_cpu_read_of_colour": {"value": px / blit_ms / 1e3, "ms_per_step": blit_ms},
               "after_cpu_rewrite_of_both_buffers": {"value": px / dirty_ms / 1e3, "ms_per_step": dirty_ms,
                                                     "note": "dirty CPU cache lines slow copy-engine writes (0.40 -> 0.58 ms at 1080p, "
                                                             "profiles/r1_e2e_ab.txt); the kernel's own stores are not affected"},
               "inout": {"value": px / inout_ms / 1e3, "ms_per_step": inout_ms, "h2d_bytes_per_step": 8 * px,
                         "d2h_bytes_per_step": 8 * px,
                         "api": "rt_render_host, in/out framebuffer cleared on the host before each call (raytracing.cpp:17,90-94); "
                                "pinned buffers: the kernel bulk-loads each tile's previous colour + depth and stores "
                                "the merged tile back"}}
        # the frame that came back must be the reference's frame (checked against the oracle, outside any timing)

    # ---- workload counters + roofline of the dominant (only) kernel
    # (N > 1: rank 0's share of the frame, the part its kernel launch actually processes)
    if orbit:
        ks = range(0, 64, max(1, 64 // min(64, args.steps)))
        cnts = [scene.render(rt.make_frame(cams[k][0], cams[k][2], c["projInv"], W, H, bands=bands),
                             want_counters=True)["counters"] for k in ks]
        cnt = {k: sum(x[k] for x in cnts) / len(cnts) for k in cnts[0]}
    else:
        r = scene.render(rt.make_frame(c["rayPos"], c["viewInv"], c["projInv"], W, H, bands=bands), want_counters=True)
        cnt = r["counters"]
    rays_launch = max(1, cnt["rays"])
    alg_bytes, alg_flops = algorithmic_work(kind, cnt, rays_launch)
    hbm, peak_kind, sm_max = peaks()
    clk = clocks.summary()
    fp32_peak = 148 * 128 * 2 * (clk["sm_mhz"] or sm_max) * 1e6 / 1e12  # TFLOP/s at the clock seen under load
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    scene_bytes = scene.info()["device_bytes"]
    traffic, traffic_src, issue = None, None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")  # per launch, from the committed ncu --set full captures
    if os.path.exists(tp):
        ent = json.load(open(tp)).get(f"{args.scene}|{args.math}")
        if ent and world == 1:
            traffic, traffic_src = ent["dram_bytes_per_launch"], ent["source"]
            # the operative bound of these kernels: warp instructions issued vs 1 per SM sub-partition per clock
            peak_issue = 148 * 4 * peaks()[2] * 1e6 / 1e9
            ach_issue = ent["warp_inst_per_launch"] / (kernel_ms * 1e-3) / 1e9
            issue = {"warp_inst_per_launch": ent["warp_inst_per_launch"], "achieved_ginst_s": ach_issue,
                     "peak_ginst_s": peak_issue, "frac": ach_issue / peak_issue, "lanes_active_of_32": ent["lanes_active"],
                     "note": "148 SMs x 4 schedulers x 1 warp instruction per clock at the max SM clock; instruction "
                             "count from the ncu capture, duration measured live"}
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                "traffic": traffic, "traffic_source": traffic_src, "issue": issue, "peak_source": f"{peak_kind} copy bandwidth (MEASURED_PEAKS.json)",
                "kernel_ms": kernel_ms, "kernel_ms_warm_l2": kernel_ms_warm,
                "algorithmic_bytes_per_launch": alg_bytes, "algorithmic_flops_per_launch": alg_flops,
                "rays_per_launch": rays_launch,
                "fp32": {"achieved_tflops": alg_flops / (kernel_ms * 1e-3) / 1e12, "peak_tflops": fp32_peak,
                         "frac": alg_flops / (kernel_ms * 1e-3) / 1e12 / fp32_peak},
                "note": ("scene fits in L2: the operative bounds are cached-gather latency and instruction issue, see DESIGN.md"
                         if scene_bytes < (100 << 20) else
                         "scene exceeds L2: gathers miss to HBM in 32-byte sectors, see DESIGN.md")}
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "reference scene file (fixture copy), viewer start camera",
        "config": {"workload": args.scene, "scene": scene_name, "width": W, "height": H, "shading": "Normal",
                   "math": args.math, "sched": args.sched, "grid_layout": args.grid_layout,
                   "scene_bytes": scene.info()["device_bytes"], "scene_build_s": round(t_build, 3),
                   "staged_peer_store_launches": staged,
                   "mesh_accel": args.mesh_accel, "mesh": scene.mesh_stats() if kind == "mesh" else None, "l2": f"flushed between steps ({L2_FLUSH_BYTES >> 20} MiB write)",
                   "partition": "whole frame" if world == 1 else
                   f"{band_rows}-row bands interleaved over {world} GPUs, " +
                   ("tiles stored by the render kernel into rank 0's framebuffer over NVLink (CUDA IPC peer mapping)"
                    if shared is not None else "NCCL send/recv gather to rank 0")},
        "clocks": clk, "gpu_launches": args.steps, "roofline": roofline,
        "counters_per_ray": {k: v / rays_launch for k, v in cnt.items()},
    }
    if multi_frame is not None:
        want = oracle_frame(g, scene_name, c, W, H)
        if want is not None:
            d = util.compare_frames(multi_frame[0], multi_frame[1], want[0], want[1])
            d["bit_identical"] = bool(util.same_bits(multi_frame[1], want[1]) and np.array_equal(multi_frame[0], want[0]))
            d["checker"] = want[2]
            line["parity"] = d
    if e2e is not None:
        line["e2e"] = e2e
        line["cpu_baseline"] = cpu_baseline(g, scene_name, c, W, H, orbit=orbit)
        # correctness of what was timed: the e2e frame against the oracle port (bit-exact in exact mode)
        want = oracle_frame(g, scene_name, c, W, H)
        if want is not None:
            wc, wt, who = want
            d = util.compare_frames(chk_c, chk_t, wc, wt)
            d["bit_identical"] = bool(util.same_bits(chk_t, wt) and np.array_equal(chk_c, wc))
            d["checker"] = who
            line["parity"] = d
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="cuda", choices=("cuda", "reference"))
    ap.add_argument("--scene", default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument("--math", default="exact", choices=("exact", "fast"))
    ap.add_argument("--sched", default="auto", choices=("auto", "tiles", "persistent", "compact"))
    ap.add_argument("--mesh-accel", default="auto", choices=("auto", "host", "lbvh_gpu"),
                    help="who builds the 8-wide BVH: the host (binned SAH) or the device (linear BVH)")
    ap.add_argument("--band-rows", type=int, default=8,
                    help="N > 1: rows per interleaved band (a multiple of the 8-row block height)")
    ap.add_argument("--grid-layout", default="auto", choices=("auto", "linear", "bricks"))
    ap.add_argument("--gather", default="p2p", choices=("p2p", "nccl"),
                    help="N > 1: how tiles reach rank 0 (fused peer stores, or an NCCL send/recv gather)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
