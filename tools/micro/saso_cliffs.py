"""SASO sketches around the streaming kernel's envelope (d = 2048, m = 2^18, n = 2048, double): aligned / misaligned leading
dimension, vec_nnz 8 / 40, both layouts, and a LASO: launches and time"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
d, m, n = 2048, 1 << 18, 2048
dev = torch.device("cuda", 0)
buf = torch.empty((m + 8) * (n + 8), dtype=torch.float64, device=dev)
rb.fill_dense(rb.DenseDist(m + 8, n + 8), buf, rb.RNGState(1))
B = torch.zeros(d * n, dtype=torch.float64, device=dev)
for name, lay, lda, k, ma in (("aligned C", "C", m, 8, "S"), ("lda odd C", "C", m + 1, 8, "S"), ("aligned R", "R", n, 8, "S"), ("lda odd R", "R", n + 1, 8, "S"),
                              ("vec_nnz 40 C", "C", m, 40, "S"), ("LASO 8 C", "C", m, 8, "L"), ("LASO 512 C", "C", m, 512, "L")):
    S = rb.SparseSkOp(rb.SparseDist(d, m, k, ma), 7)
    ldb = d if lay == "C" else n
    f = lambda: rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, buf, lda, 0.0, B, ldb)
    l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"{name:14s} launches {nl:2d} {sorted(ts)[1]:8.3f} ms", flush=True)
