#!/usr/bin/env python
"""One line per gpurun_out/bench_*.json: the numbers worth reading side by side."""
import glob, json, os, sys
d = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
print(f"{'file':34s} {'N':>2s} {'Mrays/s':>9s} {'ms':>8s} {'k_ms':>8s} {'e2e':>8s} {'e2e_ms':>7s} {'fused':>7s} {'pageabl':>7s} {'cpu':>7s} "
      f"{'cpu1':>6s} {'bound':>10s} {'frac':>6s} {'issue':>6s} {'fp32':>6s} {'gather':>6s} {'hbm':>6s} {'bitexact':>8s} {'fastout':>12s} clocks")
for f in sorted(glob.glob(os.path.join(d, "bench_*.json"))):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        err = f.replace(".json", ".err")
        print(f"{os.path.basename(f):34s} ERROR {e}: {open(err).read()[-300:] if os.path.exists(err) else ''}")
        continue
    r = j.get("roofline") or {}
    e = j.get("e2e") or {}
    p = j.get("parity") or {}
    c = j.get("cpu_baseline") or {}
    cand = r.get("candidates") or {}
    fr = lambda k: (cand.get(k) or {}).get("frac", 0)  # noqa: E731
    fm = p.get("fast_math") or {}
    fast = f"{fm.get('mask_mismatch', '-')}/{fm.get('depth_outliers', '-')}/{fm.get('rgb_outliers', '-')}"
    print(f"{os.path.basename(f)[6:-5]:34s} {j.get('n_gpus', 1):2d} {j['value']:9.0f} {j['ms_per_step']:8.4f} {r.get('kernel_ms', 0):8.4f} "
          f"{e.get('value', 0):8.0f} {e.get('ms_per_step', 0):7.3f} {(e.get('fused_clear') or e.get('inout') or {}).get('ms_per_step', 0):7.3f} "
          f"{(e.get('pageable') or {}).get('ms_per_step', 0):7.3f} {c.get('value', 0):7.1f} "
          f"{(c.get('single_thread') or {}).get('value', 0):6.2f} {str(r.get('bound')):>10s} {r.get('frac', 0):6.3f} "
          f"{(r.get('issue') or {}).get('frac', 0):6.3f} {fr('fp32'):6.3f} {fr('gather'):6.3f} {fr('hbm'):6.3f} "
          f"{str(p.get('bit_identical')):>8s} {fast:>12s} {(j.get('clocks') or {}).get('sm_mhz')}/{(j.get('clocks') or {}).get('samples')}")
