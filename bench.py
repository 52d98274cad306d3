#!/usr/bin/env python
"""bench.py — Mrays/s of the primary-ray hot path on B200, next to the reference's CPU renderer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--scene NAME] [--math exact|fast]

A "step" is one frame: every primary ray of the workload goes through the CUDA path once (fused clear + render, and for
N > 1 the gather of the screen-space bands into rank 0's framebuffer over NVLink, inside the render kernels).
The default workload (every N) is BASELINE.json configs[2], the largest single-GPU configuration and the one it names
for 1/2/4/8 GPUs: the large dense SDF grid (stand-in: example_grid.grid refined to 513^3, 540 MB, HBM-resident)
sphere-traced at 3840x2160, viewer start camera, ShadingMode::Normal, model only.  Prints ONE JSON line:

  value         device-timed whole-job Mrays/s (CUDA events around each step, L2 flushed between steps, max over ranks)
  e2e           the same frame through ONE rt_render_host call (the reference's Renderer::draw contract) with HOST
                buffers, wall clock, at N GPUs: rank 0 drives all N devices from one process (rt_init), each GPU storing
                its bands into the pinned host framebuffer over its own PCIe link
  roofline      the binding roof of the dominant kernel from MEASURED peaks (tools/peaks.cu on this box): FP32 issue,
                cached / HBM gather bandwidth, DRAM traffic; plus the instruction-issue fraction
  cpu_baseline  the reference's own C++ (oracle/_ref) on this box's host cores, all cores and one
See DESIGN.md "Measurement".
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

G65 = "example_grid|default|160x90"
WORKLOADS = {
    # name: (scene fixture, golden case holding the camera matrices for this aspect ratio, W, H, look)
    "grid65_1080p": ("example_grid", G65, 1920, 1080, "normal"),                 # configs[1]
    "cube_512": ("cube", "cube|default|128x128", 512, 512, "normal"),            # configs[0]
    "octree6_4k": ("sdf_6", "sdf_6|default|160x90", 3840, 2160, "normal"),
    "bunny_4k": ("stanford-bunny", "stanford-bunny|default|160x90", 3840, 2160, "normal"),
    "grid65_4k": ("example_grid", G65, 3840, 2160, "normal"),
    # deterministic stand-ins for the resources the reference tree does not ship (standins.py)
    "grid257_4k": ("example_grid_x4", G65, 3840, 2160, "normal"),   # configs[2], 68 MB: L2-resident
    "grid513_4k": ("example_grid_x8", G65, 3840, 2160, "normal"),   # configs[2], 540 MB: HBM-resident (DEFAULT)
    "octree8_4k": ("octree_d8", G65, 3840, 2160, "normal"),         # configs[3]
    "octree9_4k": ("octree_d9", G65, 3840, 2160, "normal"),         # configs[3]
    "bunny_s2_4k": ("stanford-bunny_s2", G65, 3840, 2160, "normal"),  # configs[4] mesh, 1.1 M triangles
    # configs[4]: step k renders frame k % 64 of the orbit (Camera::rotate(2*pi/64/0.01, 0) k times, camera.cpp:5-20)
    "bunny_s2_orbit64_4k": ("stanford-bunny_s2", G65, 3840, 2160, "normal"),
    # the viewer's default look (main.cpp:62-64,189-190): Lambert + shadow rays + one reflection bounce + ground plane
    "default_look_1080p": ("example_grid", G65, 1920, 1080, "default"),
    "default_look_octree_1080p": ("sdf_6", "sdf_6|default|160x90", 1920, 1080, "default"),
    "default_look_bunny_1080p": ("stanford-bunny", "stanford-bunny|default|160x90", 1920, 1080, "default"),
}
ORBIT_DX = float(np.float32(2.0 * np.pi / 64.0 / 0.01))
HEADLINE = "grid513_4k"
METRIC = "Mrays/s primary rays"
L2_FLUSH_BYTES = 256 << 20


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured copy bandwidth (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def measured_roofs():
    """FP32 (with / without FMA), issue rate, cached-gather rates and dependent-load latencies measured on THIS box by
    tools/peaks (hand-written micro-benchmarks), cached per box; the committed capture of an earlier box otherwise."""
    cache = os.environ.get("RT_PEAKS_CACHE", "/tmp/rt_b200_peaks.json")
    if os.path.exists(cache):
        return json.load(open(cache)), "tools/peaks on this box"
    exe = os.path.join(ROOT, "tools", "peaks")
    if os.path.exists(exe):
        try:
            r = subprocess.run([exe], capture_output=True, text=True, timeout=180)
            if r.returncode == 0:
                d = json.loads(r.stdout.strip().splitlines()[-1])
                json.dump(d, open(cache, "w"))
                return d, "tools/peaks on this box"
            sys.stderr.write("tools/peaks failed: " + r.stderr[-500:] + "\n")
        except Exception as e:  # noqa: BLE001
            sys.stderr.write(f"tools/peaks failed: {e}\n")
    p = os.path.join(ROOT, "profiles", "r2_peaks.json")
    if os.path.exists(p):
        return json.load(open(p)), "profiles/r2_peaks.json (tools/peaks on an earlier box of this pool)"
    return None, None


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
    scene_name, case_key, W, H, look = WORKLOADS[name]
    c = g.case(case_key)
    assert abs(W / H - c["W"] / c["H"]) < 1e-6
    return g, scene_name, c, W, H, look


def look_kwargs(look, plane_y):
    """Renderer's public knobs for the look (raytracing.hpp:103-106; main.cpp:62-64,189-190)."""
    if look == "default":
        return dict(shadows=True, reflections=True), plane_y
    return dict(shadows=False, reflections=False), None


def model_plane_y(arrays):
    """main.cpp:176-189: the ground plane sits at the model box's lowest y (meshes) or at -1 (SDF scenes)."""
    if "vpos4" in arrays:
        v = arrays["vpos4"]
        used = v[arrays["indices"]]
        return float((used[:, 1] / used[:, 3]).astype(np.float32).min())
    return -1.0


def ref_scene(o, arrays):
    return (o.grid(arrays["values"]) if "values" in arrays else o.octree(arrays["nodes36"]) if "nodes36" in arrays
            else o.mesh(arrays["vpos4"], arrays["indices"]))


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path (oracle/_ref when it was compiled, else the C port), all host
    threads, on the same workload: the stock Renderer::draw on the full frame when that fits the time budget
    (src/raytracing.cpp:67-102), else its pixel loop on a bounded band of rows."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    g, scene_name, c, W, H, look = load_workload(args.scene)
    from oracle import ref_oracle as R
    cores = os.cpu_count() or 1
    a = g.scene_arrays(scene_name)
    orbit = [0]
    budget_ms = 90_000.0 / max(1, args.steps + args.warmup)
    if R.available(contract=True):
        o = R.RefOracle(contract=True)  # GCC's default contraction = the reference's Release build
        o.set_threads(cores)
        s = ref_scene(o, a)
        kind = "reference"
        lk, plane_y = look_kwargs(look, model_plane_y(a))
        sd = s.with_plane(plane_y) if plane_y is not None else s
        mode = R.MODE_LAMBERT if look == "default" else R.MODE_NORMAL

        def cam():
            k = orbit[0] % 64 if "orbit" in args.scene else 0
            orbit[0] += 1
            return R.CamSpec(n_rotate=k, dx=ORBIT_DX, dy=0.0) if k else R.CamSpec()

        def full():  # the stock Renderer::draw, timed by its own stopwatch (raytracing.cpp:76,97-101)
            return sd.render_timed(cam(), W, H, frames=1, proj_inv=c["projInv"], mode=mode, **lk)[0]

        def rows(y0, y1):
            return s.render_rows(cam(), W, H, y0, y1, proj_inv=c["projInv"])[2]
        probe_rows = max(8, H // 16)
        y0 = (H - probe_rows) // 2
        ms_probe = rows(y0, y0 + probe_rows)
        whole = look == "default" or ms_probe * H / probe_rows <= budget_ms  # centre rows are the dense ones: an upper bound
    else:
        from oracle import port_oracle as P  # noqa: F401
        s = util.port_scene(g, scene_name)
        kind = "port"
        whole = False

        def rows(y0, y1):
            t0 = time.perf_counter()
            s.render(c["rayPos"], c["viewInv"], c["projInv"], W, H, rows=(y0, y1))
            return (time.perf_counter() - t0) * 1e3
        probe_rows = max(8, H // 16)
        y0 = (H - probe_rows) // 2
        ms_probe = rows(y0, y0 + probe_rows)
    if whole:
        n_rows, y0 = H, 0
        step = full
        sample = f"the full {W}x{H} frame per step through the reference's stock Renderer::draw, OpenMP dynamic over rows"
    else:
        n_rows = int(min(H, max(8, probe_rows * budget_ms / max(ms_probe, 1e-3))))
        y0 = (H - n_rows) // 2

        def step():
            return rows(y0, y0 + n_rows)
        sample = (f"rows {y0}..{y0 + n_rows} of {H} ({n_rows * W} rays per step), Renderer::draw's pixel loop on that band, "
                  f"OpenMP dynamic")
    for _ in range(args.warmup):
        step()
    total = 0.0
    for _ in range(args.steps):
        total += step()
    ms = total / args.steps
    value = n_rows * W / ms / 1e3
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "reference scene file (fixture copy) or its documented stand-in, viewer start camera",
        "config": {"workload": args.scene, "scene": scene_name, "width": W, "height": H,
                   "shading": "Lambert + shadows + reflections + plane" if look == "default" else "Normal"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ CUDA arm
def cpu_baseline(g, scene_name, c, W, H, look, orbit=False):
    """Rank 0: the reference's CPU path timed on this box's host cores on a bounded sample, all cores and one."""
    from oracle import ref_oracle as R
    cores = os.cpu_count() or 1
    try:
        if R.available(contract=True):
            o = R.RefOracle(contract=True)
            a = g.scene_arrays(scene_name)
            s = ref_scene(o, a)
            lk, plane_y = look_kwargs(look, model_plane_y(a))
            sd = s.with_plane(plane_y) if plane_y is not None else s
            mode = R.MODE_LAMBERT if look == "default" else R.MODE_NORMAL
            o.set_threads(cores)
            if orbit:
                ks = (0, 16, 32, 48)
                ms = [sd.render_timed(R.CamSpec(n_rotate=k, dx=ORBIT_DX, dy=0.0), W, H, frames=2, proj_inv=c["projInv"])[0]
                      for k in ks]
                best = float(np.mean(ms))
                sample = (f"orbit frames {ks} of 64 at {W}x{H} through the reference's Renderer::draw (best of 2 each, mean "
                          f"over the 4 cameras), scalar ISPC stand-in")
            else:
                best, _each = sd.render_timed(R.CamSpec(), W, H, frames=3, proj_inv=c["projInv"], mode=mode, **lk)
                sample = f"3 full {W}x{H} frames through the reference's Renderer::draw (best), scalar ISPC stand-in"
            # one thread: a centred band of rows sized for ~3 s, through the same pixel loop
            o.set_threads(1)
            band = max(8, H // 64)
            y0 = (H - band) // 2
            ms1 = s.render_rows(R.CamSpec(), W, H, y0, y0 + band, proj_inv=c["projInv"])[2]
            band = int(min(H, max(band, band * 3000.0 / max(ms1, 1e-3))))
            y0 = (H - band) // 2
            ms1 = s.render_rows(R.CamSpec(), W, H, y0, y0 + band, proj_inv=c["projInv"])[2]
            o.set_threads(cores)
            return {"value": W * H / best / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample,
                    "single_thread": {"value": band * W / ms1 / 1e3, "unit": "Mrays/s", "cores": 1,
                                      "sample": f"rows {y0}..{y0 + band} of {H} (the model's rows: denser than the frame "
                                                f"average), ShadingMode::Normal, 1 OpenMP thread"}}
    except Exception as e:  # noqa: BLE001 - fall through to the port
        sys.stderr.write(f"reference oracle unavailable ({e}); timing the C port\n")
    s = util.port_scene(g, scene_name)
    t0 = time.perf_counter()
    s.render(c["rayPos"], c["viewInv"], c["projInv"], W, H)
    ms = (time.perf_counter() - t0) * 1e3
    return {"value": W * H / ms / 1e3, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": f"1 full {W}x{H} frame through oracle/port_oracle.c (OpenMP)"}


def oracle_frame(g, scene_name, c, W, H, look):
    """The checker's frame for the timed workload (outside any timing). SDF scenes: the plain-C oracle. Meshes: the
    reference's own code when oracle/_ref is present (it builds its BVH); the C oracle's every-triangle loop would need
    rays x triangles tests and is only used for small meshes."""
    a = g.scene_arrays(scene_name)
    lk, plane_y = look_kwargs(look, model_plane_y(a))
    if "vpos4" not in a or a["indices"].size // 3 <= 64:
        from oracle import port_oracle as P
        ps = util.port_scene(g, scene_name)
        if plane_y is not None:
            ps = ps.with_plane(plane_y)
        wc, wt, _, _ = ps.render(c["rayPos"], c["viewInv"], c["projInv"], W, H,
                                 mode=P.MODE_LAMBERT if look == "default" else P.MODE_NORMAL, **lk)
        return wc, wt, "oracle/port_oracle.c"
    from oracle import ref_oracle as R
    if not R.available():
        return None
    o = R.RefOracle()
    s = o.mesh(a["vpos4"], a["indices"])
    if plane_y is not None:
        s = s.with_plane(plane_y)
    wc, wt, _ = s.render(R.CamSpec(), W, H, proj_inv=c["projInv"], mode=R.MODE_LAMBERT if look == "default" else R.MODE_NORMAL,
                         **lk)
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


def build_roofline(kind, scene_bytes, alg_bytes, alg_flops, kernel_ms, kernel_ms_warm, rays_launch, math, traffic_ent):
    """frac = max(flops / FP32 peak, gather bytes / gather peak, DRAM bytes / HBM peak), every peak measured."""
    hbm, hbm_src, sm_max = hbm_peak()
    roofs, roofs_src = measured_roofs()
    sec = kernel_ms * 1e-3
    cands = {}
    if roofs:
        fpeak = roofs["fp32_ffma_tflops"] if math == "fast" else roofs["fp32_nofma_tflops"]
        cands["fp32"] = {"achieved": alg_flops / sec / 1e12, "peak": fpeak, "unit": "TFLOP/s",
                         "peak_is": ("independent FFMA streams" if math == "fast" else
                                     "independent FMUL+FADD streams: the exact (no-FMA) flavour's ceiling; with FMA "
                                     f"{roofs['fp32_ffma_tflops']:.1f}")}
        # which cache level serves the gathers is decided by the scene's footprint; grids are read as 4-byte scalars,
        # octree values / BVH node runs / triangles as 16-byte vectors
        regime = ("footprint_1MB" if scene_bytes <= (4 << 20) else "l2_64MB" if scene_bytes <= (100 << 20) else "hbm_4GB")
        table = roofs["gather_useful_gbs"] if kind == "grid" else roofs["gather16_useful_gbs"]
        cands["gather"] = {"achieved": alg_bytes / sec / 1e9, "peak": table[regime], "unit": "GB/s",
                           "peak_is": f"random {'4' if kind == 'grid' else '16'}-byte gathers over a {regime} footprint "
                                      f"(useful bytes; coherent 8x4 ray tiles share lines, so this roof can be beaten)"}
    traffic = traffic_src = issue = None
    if traffic_ent:
        traffic, traffic_src = traffic_ent["dram_bytes_per_launch"], traffic_ent["source"]
        cands["hbm"] = {"achieved": traffic / sec / 1e9, "peak": hbm, "unit": "GB/s",
                        "peak_is": hbm_src + "; achieved = DRAM bytes of the ncu capture / live duration"}
        peak_issue = roofs["issue_ginst_s"] if roofs else 148 * 4 * sm_max * 1e6 / 1e9
        ach_issue = traffic_ent["warp_inst_per_launch"] / sec / 1e9
        issue = {"warp_inst_per_launch": traffic_ent["warp_inst_per_launch"], "achieved_ginst_s": ach_issue,
                 "peak_ginst_s": peak_issue, "frac": ach_issue / peak_issue, "lanes_active_of_32": traffic_ent["lanes_active"],
                 "note": "warp instructions of the ncu capture / live duration against the measured issue rate "
                         "(1 per SM sub-partition per clock)"}
    for v in cands.values():
        v["frac"] = v["achieved"] / v["peak"]
    # the binding roof: the largest fraction among those that are roofs at all for this access pattern (a gather figure
    # above 1 only says coherent loads beat random ones; it is reported, not chosen)
    usable = {k: v for k, v in cands.items() if v["frac"] <= 1.0}
    if usable:
        bound = max(usable, key=lambda k: usable[k]["frac"])
        top = usable[bound]
        out = {"bound": {"fp32": "fp32", "gather": "l2_gather" if scene_bytes <= (100 << 20) else "hbm_gather",
                         "hbm": "hbm"}[bound],
               "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"], "frac": top["frac"]}
    else:
        ach = alg_bytes / sec / 1e9
        out = {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm}
    out.update({
        "traffic": traffic, "traffic_source": traffic_src, "candidates": cands, "issue": issue,
        "peak_source": roofs_src or hbm_src, "measured_peaks": roofs,
        "hbm_accounting": {"achieved": alg_bytes / sec / 1e9, "peak": hbm, "unit": "GB/s",
                           "frac": alg_bytes / sec / 1e9 / hbm,
                           "note": "algorithmic gather bytes against the HBM copy bandwidth: an accounting figure, the "
                                   "gathers are served by L1 / L2 (see traffic)"},
        "kernel_ms": kernel_ms, "kernel_ms_warm_l2": kernel_ms_warm, "algorithmic_bytes_per_launch": alg_bytes,
        "algorithmic_flops_per_launch": alg_flops, "rays_per_launch": rays_launch})
    return out


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
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        cpu_group = dist.new_group(backend="gloo")  # host-side waits that must not occupy the GPUs
    rt.set_math_mode(rt.MATH_FAST if args.math == "fast" else rt.MATH_EXACT)
    rt.set_sched_mode({"auto": rt.SCHED_AUTO, "tiles": rt.SCHED_TILES, "persistent": rt.SCHED_PERSISTENT, "compact": rt.SCHED_COMPACT}[args.sched])

    g, scene_name, c, W, H, look = load_workload(args.scene)
    layout_kw = {}
    arrays = g.scene_arrays(scene_name)  # (stand-ins are generated here, outside the build stopwatch)
    lk, plane_y = look_kwargs(look, model_plane_y(arrays))
    mode = rt.SHADE_LAMBERT if look == "default" else rt.SHADE_NORMAL
    t_build = time.perf_counter()
    if "values" in arrays and args.grid_layout != "auto":
        layout_kw = {"layout": {"linear": rt.GRID_LINEAR, "bricks": rt.GRID_BRICKS}[args.grid_layout]}
    if "vpos4" in arrays and args.mesh_accel != "auto":
        layout_kw = {"accel": {"host": rt.MESH_BVH8, "lbvh_gpu": rt.MESH_LBVH_GPU, "brute": rt.MESH_BRUTE}[args.mesh_accel]}
    scene = util.cuda_scene(g, scene_name, **layout_kw)
    t_build = time.perf_counter() - t_build
    kind = {0: "grid", 1: "octree", 2: "mesh"}[scene.kind]
    band_rows = args.band_rows
    bands = (band_rows, world, rank) if world > 1 else None

    def mk(pos, vinv, **kw):
        return rt.make_frame(pos, vinv, c["projInv"], W, H, mode=mode, plane_y=plane_y, **lk, **kw)
    # FrameBuffer::clear (raytracing.hpp:16-19) is fused into the render launch: one kernel writes every pixel once
    frame = mk(c["rayPos"], c["viewInv"], bands=bands, clear=True)
    frames = [frame]
    orbit = "orbit" in args.scene
    cams = None
    if orbit:  # the 64 cameras of the orbit, from the host mirror of the reference's Camera (camera.cpp:5-20)
        cams = [rt.host_camera(n_rotate=k, dx=ORBIT_DX, dy=0.0) for k in range(64)]
        frames = [mk(pos, vinv, bands=bands, clear=True) for pos, _v, vinv in cams]
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
    launches = [0]

    def gather_tiles():
        """N > 1, --gather nccl: every rank ships its interleaved row bands to rank 0 (NCCL send/recv over NVLink)."""
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
        launches[0] += 1
        gather_tiles()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    step_no[0] = 0
    launches[0] = 0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    clocks = ClockSampler(local)
    clocks.__enter__()  # sampled across the device-timed loops below and a sustained loop of the same launch
    barrier()
    for a, b in ev:
        flush.zero_()  # evict the scene and the framebuffer from L2 between timed iterations (untimed)
        if world > 1:
            dist.barrier()
        a.record()
        step()
        b.record()
    barrier()
    timed_launches = launches[0]
    ms_total = sum(a.elapsed_time(b) for a, b in ev)
    if world > 1:
        tt = torch.tensor([ms_total], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    value = px / ms_step / 1e3

    # warm-L2 variant (frames back to back, as a viewer would issue them), kernel only, into a LOCAL framebuffer;
    # kept going for >= 1.2 s so that the clock sampler (50 ms period) sees the same launch under sustained load
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_warm = 0
    t_end = time.perf_counter() + 1.2
    a.record()
    while n_warm < args.steps or time.perf_counter() < t_end:
        for i in range(8):
            scene.render_device(frames[(n_warm + i) % len(frames)], color.data_ptr(), None, tbuf.data_ptr(),
                                stream=stream.cuda_stream)
        n_warm += 8
        if n_warm % 64 == 0:
            torch.cuda.synchronize()  # bound the launch queue
    b.record()
    barrier()
    kernel_ms_warm = a.elapsed_time(b) / n_warm
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
    per_rank = None
    if world > 1:  # per-rank kernel time (local framebuffer): separates load imbalance from the gather's ingress
        tt = torch.zeros(world, device=dev)
        tt[rank] = kernel_ms
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        per_rank = [float(x) for x in tt.tolist()]

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
        # rank 0 now drives ALL the GPUs from its own process for the end-to-end number: wait on the host, not with a
        # GPU-side collective that would keep a kernel spinning on this rank's device
        dist.barrier(group=cpu_group)
        dist.destroy_process_group()
        return

    # ---- end to end through ONE host-buffer C-ABI call at N GPUs (pinned host framebuffer, transfers inside the timed
    #      region): rt_init lists the N devices, the scene is replicated, rt_render_host fans the frame out
    import ctypes as C
    from triangles_sdf_cpu_raytracing_b200 import _lib as L
    n_dev = min(world, torch.cuda.device_count())
    rt.init(list(range(n_dev)))
    t_rep = time.perf_counter()
    scene_n = util.cuda_scene(g, scene_name, **layout_kw) if n_dev > 1 else scene
    t_rep = time.perf_counter() - t_rep
    hc = torch.zeros(px, dtype=torch.int32).pin_memory()
    ht = torch.full((px,), float("inf"), dtype=torch.float32).pin_memory()
    dms, tms = C.c_float(), C.c_float()
    n_e2e = max(5, min(args.steps, 30))
    frame1 = mk(c["rayPos"], c["viewInv"], clear=True)  # whole frame: the library partitions it over the devices

    def time_host(fr, before, cbuf=hc, tbuf_=ht):
        """before(i): what the CPU does to the host framebuffer between two calls (untimed, like the reference's own
        stopwatch which only brackets draw)."""
        times, kms = [], []
        for i in range(n_e2e + 3):
            before()
            t0 = time.perf_counter()
            L.check(L.lib().rt_render_host(scene_n.handle, C.byref(fr), C.c_void_p(cbuf.data_ptr()),
                                           C.c_void_p(tbuf_.data_ptr()), None, C.byref(dms), C.byref(tms), None))
            if i >= 3:
                times.append((time.perf_counter() - t0) * 1e3)
                kms.append(dms.value)
        return float(np.mean(times)), float(np.mean(kms))

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
    L.check(L.lib().rt_render_host(scene_n.handle, C.byref(frame1), C.c_void_p(hc.data_ptr()),
                                   C.c_void_p(ht.data_ptr()), None, C.byref(dms), C.byref(tms), None))
    chk_c = hc.numpy().view(np.uint32).reshape(H, W).copy()
    chk_t = ht.numpy().reshape(H, W).copy()
    # (a) THE reference's contract, literally (src/main.cpp:197,203; raytracing.cpp:17,90-94): the caller clears its
    #     FrameBuffer on the host, then ONE draw call on the in/out buffers.  Timed like the reference times itself: the
    #     stopwatch brackets draw, not the clear.  The kernels bulk-load each tile's previous depth, leave tiles without a
    #     nearer hit alone, and store the others (fetching the previous colour only for partly hit tiles).
    inout = mk(c["rayPos"], c["viewInv"])
    xfer0 = scene_n.transfer_bytes()
    e2e_ms, e2e_kms = time_host(inout, host_clear)
    xfer1 = scene_n.transfer_bytes()
    xfer_calls = n_e2e + 3
    inout_chk_c = hc.numpy().view(np.uint32).reshape(H, W).copy()
    inout_chk_t = ht.numpy().reshape(H, W).copy()
    # the same with the host-side clear inside the stopwatch (what a frame of the viewer's loop costs the caller)
    t_clear = []
    for _ in range(5):
        t0 = time.perf_counter()
        host_clear()
        t_clear.append((time.perf_counter() - t0) * 1e3)
    host_clear_ms = float(np.median(t_clear))
    # (b) clear + draw as one call (rt_frame.clear = 1, an extension): nothing but the frame parameters goes up, every
    #     pixel comes back; the caller never touches its buffers
    fused_ms, fused_kms = time_host(frame1, lambda: None)
    blit_ms, _ = time_host(frame1, host_blit)
    dirty_ms, _ = time_host(frame1, host_dirty)
    # (c) an unmodified caller: pageable std::vector-like buffers (no pin()), in/out contract, cleared on the host
    pc = torch.zeros(px, dtype=torch.int32)
    pt = torch.full((px,), float("inf"), dtype=torch.float32)

    def page_clear():
        pc.zero_()
        pt.fill_(float("inf"))
    n_keep, n_e2e = n_e2e, max(3, n_e2e // 3)
    pageable_ms, _ = time_host(inout, page_clear, pc, pt)
    n_e2e = n_keep
    # bytes that cross PCIe per in/out frame: counted by the kernels themselves as they issue the bulk copies
    # (rt_scene_transfer_bytes): previous depth in for the 32 x 4-pixel tiles in which a ray hits anything (a tile is
    # first traced against a cleared depth: no hit then means no hit with any previous depth, and nothing is fetched),
    # colour + depth out for tiles with a nearer hit, previous colour in as well for partly hit ones.  Cross-checked
    # against the frame: the stored bytes must be 8 B x 128 pixels x the tiles that hold a hit.
    Hp, Wp = (H + 3) // 4 * 4, (W + 31) // 32 * 32
    hm = np.zeros((Hp, Wp), bool)
    hm[:H, :W] = np.isfinite(chk_t)
    per_tile = hm.reshape(Hp // 4, 4, Wp // 32, 32).sum(axis=(1, 3))
    touched, partial = int((per_tile > 0).sum()), int(((per_tile > 0) & (per_tile < 128)).sum())
    fetched, stored = (xfer1[0] - xfer0[0]) // xfer_calls, (xfer1[1] - xfer0[1]) // xfer_calls
    if W % 32 == 0 and H % 4 == 0 and stored != touched * 1024:
        raise SystemExit(f"transfer accounting: {stored} bytes stored per frame, the frame has {touched} tiles with a hit")
    depth_tiles = (fetched - partial * 128 * 4) // (128 * 4)
    h2d = fetched + C.sizeof(L.RtFrame) * n_dev
    d2h = stored
    pcie_peak = 56.0 * n_dev
    e2e = {"value": px / e2e_ms / 1e3, "unit": "Mrays/s", "ms_per_step": e2e_ms, "n_gpus": n_dev,
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "device_ms_slowest_gpu": e2e_kms,
           "api": "ONE rt_render_host call per frame on the caller's in/out FrameBuffer (the reference's Renderer::draw "
                  "contract: src/main.cpp:203, raytracing.cpp:17,90-94), cleared by the caller on the host before the call "
                  "as src/main.cpp:197 does and, like the reference's own stopwatch, outside the timed region; pinned buffers"
                  + (f"; rt_init lists {n_dev} devices: the scene is replicated and the call fans interleaved 8-row "
                     f"bands out over them from this one process" if n_dev > 1 else ""),
           "transfer": "every GPU's render kernel works on the pinned host framebuffer over its own PCIe link with 128-byte "
                       "bulk copies: a tile is traced against a cleared depth first; only if a ray hits is its previous "
                       "depth fetched (pixels it bounds nearer than tFar are traced again with it, as the reference would "
                       "have), colour + depth of tiles with a nearer hit go out (previous colour in only for partly hit "
                       "tiles); one launch per GPU, no device framebuffer, no copy engine; bytes counted by the kernels "
                       "(rt_scene_transfer_bytes)",
           "tiles": {"of_32x4_pixels": int(per_tile.size), "depth_fetched": int(depth_tiles), "with_a_hit": touched,
                     "partly_hit": partial},
           "link": {"bound": "pcie", "achieved": max(h2d, d2h) / (e2e_ms * 1e-3) / 1e9, "peak": pcie_peak, "unit": "GB/s",
                    "frac": max(h2d, d2h) / (e2e_ms * 1e-3) / 1e9 / pcie_peak,
                    "peak_is": f"{n_dev} x 56 GB/s per direction (copy-engine rate of one Gen5 x16 link, tools/pcie_probe.py); "
                               f"achieved = the busier direction's bytes / wall time"},
           "scene_replication_s": round(t_rep, 3) if n_dev > 1 else None,
           "incl_host_clear": {"value": px / (e2e_ms + host_clear_ms) / 1e3, "ms_per_step": e2e_ms + host_clear_ms,
                               "host_clear_ms": host_clear_ms,
                               "note": "FrameBuffer::clear() on the host (one thread, 8 B/pixel) + the call above"},
           "fused_clear": {"value": px / fused_ms / 1e3, "ms_per_step": fused_ms, "device_ms_slowest_gpu": fused_kms,
                           "h2d_bytes_per_step": C.sizeof(L.RtFrame) * n_dev, "d2h_bytes_per_step": 8 * px,
                           "api": "rt_frame.clear = 1 (FrameBuffer::clear + draw in ONE call, an extension of the "
                                  "contract): the caller never touches its buffers, every pixel comes back",
                           "after_cpu_read_of_colour_ms": blit_ms, "after_cpu_rewrite_of_both_buffers_ms": dirty_ms},
           "pageable": {"value": px / pageable_ms / 1e3, "ms_per_step": pageable_ms, "h2d_bytes_per_step": 8 * px,
                        "d2h_bytes_per_step": 8 * px,
                        "api": "the same in/out call on pageable buffers (an integrator that never calls "
                               "FrameBuffer::pin()): staged copies through the driver's bounce buffers"}}

    # ---- workload counters + roofline of the dominant (only) kernel
    # (N > 1: rank 0's share of the frame, the part its kernel launch actually processes)
    if orbit:
        ks = range(0, 64, max(1, 64 // min(64, args.steps)))
        cnts = [scene.render(mk(cams[k][0], cams[k][2], bands=bands), want_counters=True)["counters"] for k in ks]
        cnt = {k: sum(x[k] for x in cnts) / len(cnts) for k in cnts[0]}
    else:
        cnt = scene.render(mk(c["rayPos"], c["viewInv"], bands=bands), want_counters=True)["counters"]
    rays_launch = max(1, cnt["rays"])
    alg_bytes, alg_flops = algorithmic_work(kind, cnt, rays_launch)
    scene_bytes = scene.info()["device_bytes"]
    ent = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")  # per launch, from the committed ncu --set full captures
    if os.path.exists(tp) and world == 1:
        ent = json.load(open(tp)).get(f"{args.scene}|{args.math}")
    roofline = build_roofline(kind, scene_bytes, alg_bytes, alg_flops, kernel_ms, kernel_ms_warm, rays_launch, args.math, ent)
    roofline["note"] = ("scene fits in L2: gathers are served by L1 / L2, the operative bounds are instruction issue and "
                        "dependent-load latency (measured_peaks.dependent_load_ns), see DESIGN.md"
                        if scene_bytes < (100 << 20) else
                        "scene exceeds L2: part of the gathers miss to HBM in 32-byte sectors, see DESIGN.md")
    line = {
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32",
        "data": "reference scene file (fixture copy) or its documented stand-in (standins.py), viewer start camera",
        "config": {"workload": args.scene, "scene": scene_name, "width": W, "height": H,
                   "shading": "Lambert + shadows + reflections + plane" if look == "default" else "Normal",
                   "math": args.math, "sched": args.sched, "grid_layout": args.grid_layout,
                   "scene_bytes": scene_bytes, "scene_build_s": round(t_build, 3),
                   "staged_peer_store_launches": staged,
                   "mesh_accel": args.mesh_accel, "mesh": scene.mesh_stats() if kind == "mesh" else None,
                   "l2": f"flushed between steps ({L2_FLUSH_BYTES >> 20} MiB write)",
                   "partition": "whole frame" if world == 1 else
                   f"{band_rows}-row bands interleaved over {world} GPUs, " +
                   ("tiles stored by the render kernel into rank 0's framebuffer over NVLink (CUDA IPC peer mapping)"
                    if shared is not None else "NCCL send/recv gather to rank 0")},
        "clocks": clocks.summary(), "gpu_launches": timed_launches, "roofline": roofline,
        "counters_per_ray": {k: v / rays_launch for k, v in cnt.items()},
    }
    if per_rank is not None:
        line["per_rank_kernel_ms"] = {"min": min(per_rank), "max": max(per_rank), "all": per_rank,
                                      "note": "each rank's own bands into a local framebuffer (no gather): max - min is "
                                              "load imbalance, ms_per_step - max is the gather's cost"}
    want = oracle_frame(g, scene_name, c, W, H, look)
    if multi_frame is not None and want is not None:
        d = util.compare_frames(multi_frame[0], multi_frame[1], want[0], want[1])
        d["bit_identical"] = bool(util.same_bits(multi_frame[1], want[1]) and np.array_equal(multi_frame[0], want[0]))
        d["checker"] = want[2]
        line["parity_gathered_frame"] = d
    line["e2e"] = e2e
    line["cpu_baseline"] = cpu_baseline(g, scene_name, c, W, H, look, orbit=orbit)
    if want is not None:
        # correctness of what was timed: the e2e frame against the oracle (bit-exact in exact mode)
        wc, wt, who = want
        d = util.compare_frames(inout_chk_c, inout_chk_t, wc, wt)
        d["bit_identical"] = bool(util.same_bits(inout_chk_t, wt) and np.array_equal(inout_chk_c, wc) and
                                  util.same_bits(chk_t, wt) and np.array_equal(chk_c, wc))
        d["frames_checked"] = "the in/out frame of the timed e2e call and the fused-clear frame (dirty buffers overwritten)"
        d["checker"] = who
        # the FMA-contracted flavour on the same frame: north-star tolerances, outliers counted (not timed)
        rt.set_math_mode(rt.MATH_FAST)
        L.check(L.lib().rt_render_host(scene_n.handle, C.byref(frame1), C.c_void_p(hc.data_ptr()),
                                       C.c_void_p(ht.data_ptr()), None, C.byref(dms), C.byref(tms), None))
        rt.set_math_mode(rt.MATH_FAST if args.math == "fast" else rt.MATH_EXACT)
        f = util.compare_frames(hc.numpy().view(np.uint32).reshape(H, W), ht.numpy().reshape(H, W), wc, wt)
        f["tolerances"] = "mask exact bar counted pixels, depth 1e-4 relative, RGB within 1/255"
        d["fast_math"] = f
        line["parity"] = d
    print(json.dumps(line))
    if world > 1:
        dist.barrier(group=cpu_group)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="cuda", choices=("cuda", "reference"))
    ap.add_argument("--scene", default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument("--math", default="exact", choices=("exact", "fast"))
    ap.add_argument("--sched", default="auto", choices=("auto", "tiles", "persistent", "compact"))
    ap.add_argument("--mesh-accel", default="auto", choices=("auto", "host", "lbvh_gpu", "brute"),
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
