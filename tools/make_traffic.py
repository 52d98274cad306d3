#!/usr/bin/env python
"""profiles/r<round>_*_full.txt (ncu --set full summaries; a later round's capture of a workload replaces the earlier
one) -> profiles/traffic.json: per workload the DRAM bytes and the
warp instructions of ONE launch of the render kernel, which bench.py folds into its roofline object."""
import glob, json, os, re
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r[0-9]_*_full.txt"))):
    name = os.path.basename(f)[3:-9]
    if os.path.basename(f) in ("r1_cube_512_full.txt", "r1_octree6_4k_full.txt"):  # kernels changed since (octree: round 2)
        continue  # round 1 ran cube.obj through the small-mesh kernel; RT_MESH_AUTO now builds a tree for 12 triangles
    txt = open(f).read().split("## launch 1")[0]
    def metric(m):
        r = re.search(rf"^{re.escape(m)}\s+([0-9.]+)\s*(\S*)", txt, re.M)
        return (float(r.group(1)), r.group(2)) if r else (None, None)
    rd, ru = metric("dram__bytes_read.sum")
    wr, wu = metric("dram__bytes_write.sum")
    inst, _ = metric("smsp__inst_executed.sum")
    lanes, _ = metric("smsp__thread_inst_executed_per_inst_executed.ratio")
    issue, _ = metric("smsp__issue_active.avg.pct_of_peak_sustained_active")
    dur, du = metric("gpu__time_duration.sum")
    kern = re.search(r"^## launch 0: (.*)$", txt, re.M).group(1)
    if rd is None or inst is None:
        continue
    out[f"{name}|exact"] = {
        "dram_bytes_per_launch": int(rd * UNIT[ru] + wr * UNIT[wu]), "warp_inst_per_launch": int(inst),
        "lanes_active": lanes, "issue_active_pct": issue, "ncu_duration_us": dur * (1e3 if du == "ms" else 1.0),
        "kernel": kern, "source": f"profiles/{os.path.basename(f)} (launch 0 of an `ncu --set full --clock-control none` capture)"}
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps({k: (v["dram_bytes_per_launch"], v["warp_inst_per_launch"]) for k, v in out.items()}, indent=1))
