import sys, torch
sys.path.insert(0, "/root/repo")
import randblas_b200 as rb
n = 32768
buf = torch.empty(n * n, dtype=torch.float32, device="cuda")
for fam in ("U", "G"):
    f = lambda: rb.fill_dense(rb.DenseDist(n, n, fam), buf, rb.RNGState(0))
    f(); torch.cuda.synchronize()
    ts = []
    for _ in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[3]
    print(fam, round(t, 4), "ms", round(n * n / t / 1e6, 1), "Gsamples/s")
