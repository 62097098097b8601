"""dense sketches of tall-skinny data (d = 1024 or 512, m = 65536, n = 4 .. 64), both precisions: which kernel runs and how long"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
m = 65536
dev = torch.device("cuda", 0)
for tdt, d in ((torch.float32, 1024), (torch.float64, 512)):
    for n in (4, 8, 16, 32, 48, 64):
        A = torch.empty(m * n, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
        C = torch.zeros(d * n, dtype=tdt, device=dev)
        for fam in "GU":
            S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 0, dtype=tdt)
            f = lambda: rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, C, d)
            l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            print(str(tdt)[6:], "n", n, fam, "launches", nl, round(sorted(ts)[2], 3), "ms", flush=True)
