"""python scripts/pcie_probe.py -- pinned host<->device copy rates on this box (context for bench.py's e2e line)."""
import json, time, torch
n = 2 << 30
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.uint8).pin_memory()
out = {}
for name, (dst, src) in {"d2h": (h, d), "h2d": (d, h)}.items():
    dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    out[name + "_GBps"] = round(3 * n / (time.perf_counter() - t) / 1e9, 1)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(3):
    with torch.cuda.stream(s1): h.copy_(d, non_blocking=True)
    with torch.cuda.stream(s2): d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize()
out["both_directions_each_GBps"] = round(3 * n / (time.perf_counter() - t) / 1e9, 1)
print(json.dumps(out))
