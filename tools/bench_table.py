#!/usr/bin/env python
"""One line per gpurun_out/bench_*.json: the numbers worth reading side by side."""
import glob, json, os, sys
d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
print(f"{'file':38s} {'Mrays/s':>9s} {'ms':>8s} {'k_ms':>8s} {'e2e':>8s} {'cpu':>7s} {'hbm_frac':>8s} {'bitexact':>8s} build_s clocks")
for f in sorted(glob.glob(os.path.join(d, "bench_*.json"))):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        err = f.replace(".json", ".err")
        print(f"{os.path.basename(f):38s} ERROR {e}: {open(err).read()[-300:] if os.path.exists(err) else ''}")
        continue
    r = j.get("roofline") or {}
    e = j.get("e2e") or {}
    p = j.get("parity") or {}
    c = j.get("cpu_baseline") or {}
    print(f"{os.path.basename(f)[6:-5]:38s} {j['value']:9.0f} {j['ms_per_step']:8.4f} {r.get('kernel_ms', 0):8.4f} "
          f"{e.get('value', 0):8.0f} {c.get('value', 0):7.1f} {r.get('frac', 0):8.3f} {str(p.get('bit_identical')):>8s} "
          f"{j.get('config', {}).get('scene_build_s')} {(j.get('clocks') or {}).get('sm_mhz')}")
