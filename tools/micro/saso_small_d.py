"""SASO sketches with a small sketch dimension (m = 2^18, n = 2048, vec_nnz = 8, double): the streaming kernel's records hold
three entries per row and chunk, so few rows mean long tails"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
m, n = 1 << 18, 2048
dev = torch.device("cuda", 0)
A = torch.empty(m * n, dtype=torch.float64, device=dev)
rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(1))
for d in (16, 64, 128, 256, 512, 1024, 2048, 4096):
    for lay, lda, ldb in (("C", m, d), ("R", n, n)):
        S = rb.SparseSkOp(rb.SparseDist(d, m, 8), 7)
        B = torch.zeros(d * n, dtype=torch.float64, device=dev)
        f = lambda: rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
        l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        print(f"d {d:5d} {lay} launches {nl:2d} {sorted(ts)[1]:8.3f} ms", flush=True)
