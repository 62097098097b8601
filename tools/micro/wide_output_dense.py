"""dense float sketch with a wide output (d = 2048, m = 4096, n = 65536: B is 537 MB), both layouts: how much the layout of B costs"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
for tdt, d, m, n in ((torch.float32, 2048, 4096, 65536), (torch.float32, 256, 4096, 262144), (torch.float64, 1024, 2048, 65536)):
    A = torch.empty(m * n, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    B = torch.zeros(d * n, dtype=tdt, device=dev)
    for fam in "GU":
        S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 0, dtype=tdt)
        for lay in (L.ColMajor, L.RowMajor):
            c = lay == L.ColMajor
            f = lambda: rb.sketch_general(lay, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m if c else n, 0.0, B, d if c else n)
            l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            t = sorted(ts)[2]
            print(str(tdt)[6:], "d", d, "m", m, "n", n, fam, lay.name, "launches", nl, round(t, 3), "ms", round(2.0 * d * m * n / t / 1e9, 1), "TFLOP/s", flush=True)
