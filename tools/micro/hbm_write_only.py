#!/usr/bin/env python
"""Write-only HBM ceiling on this box: cudaMemset-class fills of 2^32 floats (17.2 GB), CUDA events, median of 5.
Context for K1 (fill_dense writes and never reads): MEASURED_PEAKS.json's hbm_gbs is a COPY figure (read + write)."""
import json, statistics, torch
n = 1 << 32
x = torch.empty(n, dtype=torch.float32, device="cuda")
out = {}
for name, fn in (("zero_", lambda: x.zero_()), ("fill_", lambda: x.fill_(1.5))):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = statistics.median(ts)
    out[name] = {"ms": t, "write_GBps": 4.0 * n / t / 1e6}
print(json.dumps(out))
