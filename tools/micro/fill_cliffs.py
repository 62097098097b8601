"""fill_dense around its loops (2^28 entries): natural layout, the transposed layout, windows at aligned / unaligned offsets, float / double"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
n = 16384
dev = torch.device("cuda", 0)
L = rb.Layout
for tdt in (torch.float32, torch.float64):
    buf = torch.empty(n * n, dtype=tdt, device=dev)
    D = rb.DenseDist(n, 2 * n, "G")          # wide: natural layout RowMajor
    Dbig = rb.DenseDist(2 * n, 4 * n, "G")
    cases = [("whole, natural (RowMajor)", lambda: rb.fill_dense(rb.DenseDist(n, n, "G"), buf, rb.RNGState(0))),
             ("whole, ColMajor request", lambda: rb.fill_dense(L.ColMajor, rb.DenseDist(n, n + 4, "G"), n, n, 0, 0, buf, rb.RNGState(0))),
             ("window aligned (0, 8)", lambda: rb.fill_dense(L.RowMajor, Dbig, n, n, 0, 8, buf, rb.RNGState(0))),
             ("window unaligned (3, 5)", lambda: rb.fill_dense(L.RowMajor, Dbig, n, n, 3, 5, buf, rb.RNGState(0))),
             ("window ColMajor (3, 5)", lambda: rb.fill_dense(L.ColMajor, Dbig, n, n, 3, 5, buf, rb.RNGState(0))),
             ("uniform whole", lambda: rb.fill_dense(rb.DenseDist(n, n, "U"), buf, rb.RNGState(0)))]
    for name, f in cases:
        f(); torch.cuda.synchronize(); ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        t = sorted(ts)[1]
        print(f"{str(tdt)[6:]:8s} {name:28s} {t:8.3f} ms {n * n / t / 1e6:8.1f} Gsamples/s", flush=True)
