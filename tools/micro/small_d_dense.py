"""dense sketches with a small sketch dimension d (m = 65536, n = 1024), both precisions and layouts: which kernel runs and how long"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
m, n = 65536, 1024
dev = torch.device("cuda", 0)
for tdt in (torch.float32, torch.float64):
    A = torch.empty(m * n, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    for d in (8, 16, 32, 48, 64):
        C = torch.zeros(d * n, dtype=tdt, device=dev)
        for lay, lda, ldb in ((rb.Layout.ColMajor, m, d), (rb.Layout.RowMajor, n, n)):
            S = rb.DenseSkOp(rb.DenseDist(d, m, "G"), 0, dtype=tdt)
            f = lambda: rb.sketch_general(lay, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, C, ldb)
            l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            print(str(tdt)[6:], "d", d, lay.name, "launches", nl, round(sorted(ts)[2], 3), "ms", flush=True)
