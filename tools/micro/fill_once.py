import sys, torch
sys.path.insert(0, "/root/repo")
import randblas_b200 as rb
n = 32768
buf = torch.empty(n * n, dtype=torch.float32, device="cuda")
for fam in ("U", "G"):
    rb.fill_dense(rb.DenseDist(n, n, fam), buf, rb.RNGState(0))
torch.cuda.synchronize()
print("ok")
