"""One Gaussian and one uniform fill_dense of 2^28 floats (for ncu captures of fill_dense_kernel)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, randblas_b200 as rb
n = 16384
buf = torch.empty(n * n, dtype=torch.float32, device="cuda")
for fam in ("G", "U"):
    for _ in range(2):
        rb.fill_dense(rb.DenseDist(n, n, fam, "L"), buf, rb.RNGState(0))
torch.cuda.synchronize()
