#!/usr/bin/env python
"""Turns ncu artefacts brought back in gpurun_out/ into the small text summaries committed under profiles/.

    python tools/ncu_summary.py report  gpurun_out/prof_x.ncu-rep  > profiles/r1_x_full.txt
    python tools/ncu_summary.py launches gpurun_out/launches_x.csv > profiles/r1_x_launches.txt
"""
import csv
import subprocess
import sys
from collections import OrderedDict

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def report(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu --set full --clock-control none: {path}")
    for k, r in enumerate(data):
        print(f"\n## launch {k}: {r[col['Kernel Name']]}  grid {r[col.get('Grid Size', 0)]} block {r[col.get('Block Size', 0)]}")
        for m in METRICS:
            if m in col:
                print(f"{m:70s} {r[col[m]]:>16s} {units[col[m]]}")
        stalls = [(float(r[i]), h) for h, i in col.items()
                  if h.startswith("smsp__average_warp_latency_issue_stalled") or
                  (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"))
                  if r[i].replace(".", "").isdigit()]
        for v, h in sorted(stalls, reverse=True)[:8]:
            print(f"{h:70s} {v:16.3f}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    lines, tot = [], 0
    for r in rows[3:]:
        if len(r) > 7 and r[0].strip().isdigit() and r[7].strip().isdigit():
            n = int(r[7])
            tot += n
            lines.append((n, int(r[0]), r[1].strip()[:100]))
    if tot:
        print(f"\n## hottest source lines of the first launch (share of {tot} warp instructions executed)")
        for n, ln, text in sorted(lines, reverse=True)[:25]:
            print(f"{100.0 * n / tot:5.1f}%  L{ln}: {text}")


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit()]
    agg = OrderedDict()
    for r in rows:
        name = r[4].split("(")[0][:90]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[14]) / 1e3
    total = sum(a[1] for a in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {path}")
    print(f"# {len(rows)} launches, {total:.1f} us in total (cold-cache, serialised: compare shares, not absolutes)")
    print(f"{'kernel':92s} {'launches':>8s} {'total us':>10s} {'avg us':>9s} {'share':>7s}")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name:92s} {n:8d} {us:10.1f} {us / n:9.2f} {100 * us / total:6.1f}%")


if __name__ == "__main__":
    {"report": report, "launches": launches}[sys.argv[1]](sys.argv[2])
