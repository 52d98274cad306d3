"""CPU analysis (no GPU): how much SIMD efficiency a one-thread-per-ray kernel can reach on a workload, per pixel -> warp
mapping.  The per-ray CUDA source is run on the CPU (tests/host_emul) to get every primary ray's work items (node
visits, leaf visits, SDF samples, triangle tests); a warp executing 32 rays in lock step takes max-over-lanes steps of
each loop, so   efficiency = sum over rays of work / (32 * sum over warps of max over the warp's rays)
is the fraction of lane-slots doing useful work in that loop if nothing else interfered.  Mappings: w x h pixel tiles
per warp, and an "oracle sort" (rays sorted by their work, the bound of any compaction scheme).
usage: python tools/warp_efficiency.py [scene[:WxH] ...]   (default: the bench workloads at 1920x1080)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import triangles_sdf_cpu_raytracing_b200 as rt  # noqa: E402
from tests import host_emul as E  # noqa: E402
from tests import util  # noqa: E402

SHAPES = ((32, 1), (16, 2), (8, 4), (4, 8), (2, 16), (1, 32))
NAMES = ("inner visits", "leaf visits", "sdf samples", "tri tests")


def efficiency(work, w, h):
    H, W = work.shape
    Hp, Wp = (H + h - 1) // h * h, (W + w - 1) // w * w
    pad = np.zeros((Hp, Wp), work.dtype)
    pad[:H, :W] = work
    tiles = pad.reshape(Hp // h, h, Wp // w, w).transpose(0, 2, 1, 3).reshape(-1, w * h)
    return float(work.sum()) / max(1.0, 32.0 * float(tiles.max(axis=1).sum()))


def main():
    g = util.Golden()
    specs = sys.argv[1:] or ["example_grid:1920x1080", "sdf_6:1920x1080", "octree_d8:1920x1080", "stanford-bunny:1920x1080"]
    for spec in specs:
        name, _, size = spec.partition(":")
        W, H = (int(v) for v in (size or "1920x1080").split("x"))
        a = g.scene_arrays(name)
        s = (E.EmulScene.grid(a["values"]) if "values" in a else E.EmulScene.octree(a["nodes36"]) if "nodes36" in a
             else E.EmulScene.mesh(a["vpos4"], a["indices"]))
        pos, _view, vinv = rt.host_camera()
        pinv = rt.host_projection(W, H)[1]
        work = s.work_map(rt.make_frame(pos, vinv, pinv, W, H))
        print(f"== {name} {W}x{H}: rays {W * H}, mean work/ray " +
              ", ".join(f"{NAMES[k]} {work[..., k].mean():.2f}" for k in range(4) if work[..., k].any()))
        for k in range(4):
            wk = work[..., k].astype(np.int64)
            if not wk.any():
                continue
            row = "  ".join(f"{w}x{h} {efficiency(wk, w, h):.3f}" for w, h in SHAPES)
            srt = np.sort(wk.reshape(-1))
            n = srt.size // 32 * 32
            best = float(srt.sum()) / (32.0 * float(srt[:n].reshape(-1, 32).max(axis=1).sum() + (srt[n:].max() if n < srt.size else 0)))
            busy = float((wk > 0).mean())
            print(f"   {NAMES[k]:13s} rays with any {busy:.3f} | {row} | sorted-by-work bound {best:.3f}")


if __name__ == "__main__":
    main()
