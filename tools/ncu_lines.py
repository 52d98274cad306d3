#!/usr/bin/env python
"""Per-source-line view of an ncu report's first launch: warp instructions, share, lanes active.
   python tools/ncu_lines.py gpurun_out/prof_x.ncu-rep [min_share_pct]"""
import csv, subprocess, sys
path = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
fname, hdr, per = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Name":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = {h: i for i, h in enumerate(r)}
    elif hdr and len(r) > 10 and r[0].strip().isdigit():
        try:
            n = int(r[hdr["Instructions Executed"]]); tn = int(r[hdr["Thread Instructions Executed"]])
        except ValueError:
            continue
        if n:
            per.append((fname, int(r[0]), r[1].strip()[:110], n, tn))
tot = sum(p[3] for p in per); ttot = sum(p[4] for p in per)
print(f"# {path}: {tot} warp instructions, {ttot / tot:.2f} lanes active on average")
cum = 0
for f, ln, text, n, tn in per:
    if 100.0 * n / tot >= min_share:
        print(f"{100.0 * n / tot:5.2f}%  lanes {tn / n:5.1f}  {f}:{ln}: {text}")
