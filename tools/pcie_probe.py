"""PCIe copy bandwidth between pinned host memory and the GPU (what bounds the host-framebuffer e2e number)."""
import json, subprocess, torch
out = {}
for mb in (16, 64, 256):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for name, (dst, src) in {"d2h": (h, d), "h2d": (d, h)}.items():
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            dst.copy_(src, non_blocking=True)
        b.record()
        torch.cuda.synchronize()
        out[f"{name}_{mb}MiB_GBps"] = round(n * 10 / (a.elapsed_time(b) * 1e-3) / 1e9, 2)
try:
    q = subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current",
                        "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    out["pcie_link"] = q
except OSError:
    pass
print(json.dumps(out))
