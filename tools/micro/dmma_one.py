"""one pre-filled DMMA GEMM launch at a quarter of the C5 slab (d = 8192, m = 16384, n = 4096) for ncu"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
d, m, n = 8192, 16384, 4096
A = torch.empty(m * n, dtype=torch.float64, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
B = torch.zeros(d * n, dtype=torch.float64, device=dev)
S = rb.DenseSkOp(rb.DenseDist(d, m), 0, dtype=torch.float64); rb.fill_dense(S)
for _ in range(2):
    rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d)
torch.cuda.synchronize()
