"""SASO sketches of tall-skinny data (d = 2048, m = 2^20, vec_nnz = 8, double, n = 1 .. 128): time against the bytes of A"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
d, m = 2048, 1 << 20
dev = torch.device("cuda", 0)
for n in (1, 4, 16, 32, 64, 128):
    A = torch.empty(m * n, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(m, max(n, 1)), A, rb.RNGState(1))
    for lay, lda, ldb in (("C", m, d), ("R", n, n)):
        S = rb.SparseSkOp(rb.SparseDist(d, m, 8), 7)
        B = torch.zeros(d * n, dtype=torch.float64, device=dev)
        f = lambda: rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
        l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        t = sorted(ts)[1]
        print(f"n {n:4d} {lay} launches {nl:2d} {t:8.3f} ms  {8.0 * m * n / t / 1e6:8.1f} GB/s of A", flush=True)
