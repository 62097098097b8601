"""Right sketch B (n x d) = A (n x m) * S (m x d) with a tall SASO S: time of the streaming path (C2 shape, m = 2^18)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, randblas_b200 as rb
d, n, m = 2048, 2048, 1 << 18
A = torch.randn(n * m, dtype=torch.float64, device="cuda")      # n x m, ColMajor (lda = n): rows of A^T contiguous
B = torch.zeros(n * d, dtype=torch.float64, device="cuda")
S = rb.SparseSkOp(rb.SparseDist(m, d, 8), 1997)
def t():
    for _ in range(2): rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, n, d, m, 1.0, A, n, S, 0, 0, 0.0, B, n)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, n, d, m, 1.0, A, n, S, 0, 0, 0.0, B, n); e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)
ms = t()
print(json.dumps({"right_sketch_ms": ms, "algorithmic_GBps": (8.0 * n * m + 8.0 * n * d) / ms / 1e6}))
