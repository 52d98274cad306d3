#!/usr/bin/env python
"""BASELINE.json configs[4]: the 64-frame orbit of the large mesh over n GPUs, partitioned two ways.

    python tools/orbit_frame_parallel.py [--scene bunny_s2_orbit64_4k] [--frames 64]

  tile-parallel   every frame is split into interleaved 8-row bands over the n GPUs by ONE rt_render_device call
                  (rt_init; bands reach GPU 0's framebuffer over NVLink inside the render kernels), frames one after
                  the other: the partition BASELINE.json names;
  frame-parallel  frame k is rendered whole by GPU k % n into that GPU's own HBM (no traffic while rendering), and ONE
                  gather at the end brings every frame to GPU 0 (peer copies over NVLink).
One process drives all GPUs.  Camera k = the viewer's start camera after k calls of Camera::rotate(2*pi/64/0.01, 0)
(src/camera.cpp:5-20), from the host mirror of the reference's camera.  Frames are compared bit for bit between the two
partitions and with the n = 1 render.  Times: wall clock around the whole orbit (launch to last byte on GPU 0)."""
import argparse
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

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="bunny_s2_orbit64_4k")
ap.add_argument("--frames", type=int, default=64)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()

g, scene_name, c, W, H, look = bench.load_workload(args.scene)
px = W * H
F = args.frames
cams = [rt.host_camera(n_rotate=k, dx=bench.ORBIT_DX, dy=0.0) for k in range(F)]
frames = [rt.make_frame(pos, vinv, c["projInv"], W, H, clear=True) for pos, _v, vinv in cams]
avail = torch.cuda.device_count()
rows = []
ref_frames = None
for n in (1, 2, 4, 8):
    if n > avail:
        break
    devs = list(range(n))
    # ---- tile-parallel: one scene replicated on the n devices, one call per frame
    rt.init(devs)
    s_all = util.cuda_scene(g, scene_name)
    d0 = torch.device("cuda", 0)
    with torch.cuda.device(d0):
        out_c = torch.zeros((F, px), dtype=torch.int32, device=d0)
        out_t = torch.zeros((F, px), dtype=torch.float32, device=d0)
        st0 = torch.cuda.current_stream(d0)

        def tile_parallel():
            for k in range(F):
                s_all.render_device(frames[k], out_c[k].data_ptr(), None, out_t[k].data_ptr(), stream=st0.cuda_stream)
            st0.synchronize()
        tile_parallel()
        best_tile = 1e30
        for _ in range(args.reps):
            t0 = time.perf_counter()
            tile_parallel()
            best_tile = min(best_tile, (time.perf_counter() - t0) * 1e3)
        keep = (0, F // 3, F - 1)
        got = {k: (out_c[k].cpu().numpy().copy(), out_t[k].cpu().numpy().copy()) for k in keep}
    if ref_frames is None:
        ref_frames = got
    for k in keep:
        assert np.array_equal(got[k][0], ref_frames[k][0]) and util.same_bits(got[k][1], ref_frames[k][1]), ("tile", n, k)
    s_all.destroy()
    # ---- frame-parallel: one single-device scene per GPU, frame k on GPU k % n, one gather at the end
    scenes, bufs, streams = [], [], []
    for d in devs:
        rt.init([d])
        scenes.append(util.cuda_scene(g, scene_name))
        dev = torch.device("cuda", d)
        mine = len(range(d, F, n))
        with torch.cuda.device(dev):
            bufs.append((torch.zeros((mine, px), dtype=torch.int32, device=dev),
                         torch.zeros((mine, px), dtype=torch.float32, device=dev)))
            streams.append(torch.cuda.Stream(dev))
    with torch.cuda.device(d0):
        out_c.zero_()
        out_t.zero_()
    torch.cuda.synchronize()

    def frame_parallel(gather=True):
        for j in range((F + n - 1) // n):  # round-robin issue so every GPU starts at once
            for d in devs:
                k = j * n + d
                if k < F:
                    scenes[d].render_device(frames[k], bufs[d][0][j].data_ptr(), None, bufs[d][1][j].data_ptr(),
                                            stream=streams[d].cuda_stream)
        if gather:  # ONE gather at the end: each GPU's frames to their slots on GPU 0
            for d in devs:
                with torch.cuda.stream(streams[d]):
                    for j, k in enumerate(range(d, F, n)):
                        out_c[k].copy_(bufs[d][0][j], non_blocking=True)
                        out_t[k].copy_(bufs[d][1][j], non_blocking=True)
        for d in devs:
            streams[d].synchronize()
    frame_parallel()
    best_fp = best_fp_nogather = 1e30
    for _ in range(args.reps):
        t0 = time.perf_counter()
        frame_parallel()
        best_fp = min(best_fp, (time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter()
        frame_parallel(gather=False)
        best_fp_nogather = min(best_fp_nogather, (time.perf_counter() - t0) * 1e3)
    with torch.cuda.device(d0):
        for k in keep:
            assert np.array_equal(out_c[k].cpu().numpy(), ref_frames[k][0]), ("frame-parallel colour", n, k)
            assert util.same_bits(out_t[k].cpu().numpy(), ref_frames[k][1]), ("frame-parallel depth", n, k)
    for s in scenes:
        s.destroy()
    del bufs, out_c, out_t
    row = dict(scene=args.scene, n=n, frames=F, tile_parallel_ms=best_tile, frame_parallel_ms=best_fp,
               frame_parallel_no_gather_ms=best_fp_nogather,
               tile_parallel_mrays=F * px / best_tile / 1e3, frame_parallel_mrays=F * px / best_fp / 1e3)
    rows.append(row)
    print(json.dumps(row), flush=True)
rt.init([])
print(f"\n{'n':>2s} {'tile-parallel ms':>17s} {'frame-parallel ms':>18s} {'(no gather)':>12s} {'tile Mrays/s':>13s} {'frame Mrays/s':>14s}   ({F} frames of {W}x{H}, bit-identical)")
for r in rows:
    print(f"{r['n']:2d} {r['tile_parallel_ms']:17.2f} {r['frame_parallel_ms']:18.2f} {r['frame_parallel_no_gather_ms']:12.2f} "
          f"{r['tile_parallel_mrays']:13.0f} {r['frame_parallel_mrays']:14.0f}")
