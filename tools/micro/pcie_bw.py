"""Pinned host -> device and device -> host copy bandwidth on this box (the ceiling of bench.py's e2e leg)."""
import json, time, torch
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
out = {}
for name, (dst, src) in {"h2d": (d, h), "d2h": (h, d)}.items():
    dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); dst.copy_(src, non_blocking=True); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    out[name + "_GBps"] = n / best / 1e9
# strided 2D source like a column-major slab: 2048 rows of 256 KiB, source pitch 8 MiB
h2 = torch.empty(2048 * (8 << 20) // 8, dtype=torch.float64, pin_memory=True)
v = h2.view(2048, (8 << 20) // 8)[:, :32768]
d2 = torch.empty(2048, 32768, dtype=torch.float64, device="cuda")
d2.copy_(v, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter(); d2.copy_(v, non_blocking=True); torch.cuda.synchronize()
out["h2d_2d_GBps"] = d2.numel() * 8 / (time.perf_counter() - t0) / 1e9
print(json.dumps(out))
