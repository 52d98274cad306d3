"""bench.py's pure-host pieces (no GPU): workload table, algorithmic work, the roofline object built from the committed
measured peaks and ncu captures.  Guards the driver's bench run against a typo in a code path only the GPU box reaches."""
import json
import os

import bench


def test_workloads_resolve_and_default_is_configs_2():
    assert bench.HEADLINE == "grid513_4k" and bench.WORKLOADS[bench.HEADLINE][2:4] == (3840, 2160)
    g = None
    for name, (scene, case, W, H, look) in bench.WORKLOADS.items():
        assert look in ("normal", "default")
        if scene in ("example_grid", "sdf_6", "cube", "stanford-bunny"):
            g, scene_name, c, w, h, lk = bench.load_workload(name)
            assert (w, h) == (W, H) and c["projInv"] is not None


def test_roofline_object_from_committed_peaks_and_captures(tmp_path, monkeypatch):
    peaks = json.load(open(os.path.join(bench.ROOT, "profiles", "r2_peaks.json")))
    traffic = json.load(open(os.path.join(bench.ROOT, "profiles", "traffic.json")))
    cache = tmp_path / "peaks.json"  # what measured_roofs() caches after running tools/peaks on a box
    json.dump(peaks, open(cache, "w"))
    monkeypatch.setenv("RT_PEAKS_CACHE", str(cache))
    rays = 3840 * 2160
    cnt = dict(rays=rays, hits=int(0.24 * rays), grid_samples=int(6.49 * rays), oct_inner=0, oct_leaf=0, oct_leaf_samples=0,
               bvh_inner=0, bvh_leaf=0, bvh_tris=0)
    b, f = bench.algorithmic_work("grid", cnt, rays)
    assert abs(b / rays - (32 * 6.49 + 8)) < 1e-3 and abs(f / rays - (150 + 60 * 6.49)) < 1e-3
    r = bench.build_roofline("grid", 540 << 20, b, f, 0.366, 0.36, rays, "exact", traffic["grid513_4k|exact"])
    assert r["bound"] in ("fp32", "hbm", "l2_gather", "hbm_gather") and 0 < r["frac"] <= 1.0
    assert set(r["candidates"]) == {"fp32", "gather", "hbm"}
    assert abs(r["candidates"]["fp32"]["frac"] - f / 0.366e-3 / 1e12 / peaks["fp32_nofma_tflops"]) < 1e-9
    assert 0.5 < r["issue"]["frac"] < 1.0 and r["traffic"] == traffic["grid513_4k|exact"]["dram_bytes_per_launch"]
    # the random-gather roof is beaten by coherent tiles: reported, not chosen
    assert r["candidates"]["gather"]["frac"] > 1.0 and r["bound"] != "hbm_gather"
    # octree: 16-byte gathers over an L2-resident footprint
    cnt = dict(cnt, grid_samples=0, oct_inner=int(6.1 * rays), oct_leaf=int(5.5 * rays), oct_leaf_samples=int(2.9 * rays))
    b, f = bench.algorithmic_work("octree", cnt, rays)
    r = bench.build_roofline("octree", 22 << 20, b, f, 1.055, 1.05, rays, "exact", traffic["octree8_4k|exact"])
    assert r["candidates"]["gather"]["peak"] == peaks["gather16_useful_gbs"]["l2_64MB"] and 0 < r["frac"] <= 1.0
    # no capture for the workload: still a line, no traffic
    r = bench.build_roofline("mesh", 8 << 20, 1e9, 1e10, 0.3, 0.3, rays, "fast", None)
    assert r["traffic"] is None and r["candidates"]["fp32"]["peak"] == peaks["fp32_ffma_tflops"]
