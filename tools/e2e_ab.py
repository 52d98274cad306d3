"""A/B of rt_render_host's transfer modes (RT_HOST_ZEROCOPY 0 / 1 / 2) on pinned host framebuffers.
One subprocess per mode (the knob is read once at library load); prints mean ms per call.
usage: [E2E_SCENES=scene:WxH,...] python tools/e2e_ab.py [modes, default 0,1,2]"""
import os
import subprocess
import sys

CHILD = r"""
import ctypes as C, time, sys
import numpy as np, torch
import triangles_sdf_cpu_raytracing_b200 as rt
from triangles_sdf_cpu_raytracing_b200 import _lib as L
from tests import util
g = util.Golden()
mode = sys.argv[1]
import os
cases = [(c.split(":")[0],) + tuple(int(v) for v in c.split(":")[1].split("x"))
         for c in os.environ.get("E2E_SCENES", "example_grid:1920x1080,example_grid:3840x2160,stanford-bunny:3840x2160").split(",")]
for scene_name, W, H in cases:
    s = util.cuda_scene(g, scene_name)
    c = g.case([k for k in g.cases if k.startswith(scene_name + "|default")][0])
    pinv = rt.host_projection(W, H)[1]
    f = rt.make_frame(c["rayPos"], c["viewInv"], pinv, W, H, clear=True)
    color = torch.zeros(W * H, dtype=torch.int32).pin_memory()
    t = torch.zeros(W * H, dtype=torch.float32).pin_memory()
    dms, tms = C.c_float(), C.c_float()
    def call():
        L.check(L.lib().rt_render_host(s.handle, C.byref(f), C.c_void_p(color.data_ptr()), C.c_void_p(t.data_ptr()), None,
                                       C.byref(dms), C.byref(tms), None))
    for _ in range(5):
        call()
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); call(); ts.append((time.perf_counter() - t0) * 1e3)
    hit = int((t.numpy() < 1e30).sum())
    td = []
    for _ in range(10):
        color.fill_(0x5A5A5A5A); t.fill_(0.25)
        t0 = time.perf_counter(); call(); td.append((time.perf_counter() - t0) * 1e3)
    f = rt.make_frame(c["rayPos"], c["viewInv"], pinv, W, H)   # in/out contract: the caller clears on the host
    ti = []
    for _ in range(12):
        color.zero_(); t.fill_(float("inf"))
        t0 = time.perf_counter(); call(); ti.append((time.perf_counter() - t0) * 1e3)
    hit2 = int((t.numpy() < 1e30).sum())
    assert hit2 == hit, (hit, hit2)
    print(f"mode {mode} {scene_name:15s} {W}x{H}  mean {np.mean(ts):.4f} ms  min {np.min(ts):.4f}  after-cpu-rewrite {np.mean(td):.4f}  inout {np.mean(ti[2:]):.4f}  "
          f"floor56 {W*H*8/56e9*1e3:.4f}  hits {hit}  csum {int(color.numpy().view(np.uint32).astype(np.uint64).sum())}", flush=True)
"""

if __name__ == "__main__":
    modes = (sys.argv[1] if len(sys.argv) > 1 else "0,1,2").split(",")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for m in modes:
        env = dict(os.environ, RT_HOST_ZEROCOPY=m)
        r = subprocess.run([sys.executable, "-c", CHILD, m], env=env, cwd=root, capture_output=True, text=True, timeout=300)
        sys.stdout.write(r.stdout)
        if r.returncode != 0:
            sys.stdout.write(f"mode {m} FAILED rc={r.returncode}\n{r.stderr[-1500:]}\n")
        sys.stdout.flush()
