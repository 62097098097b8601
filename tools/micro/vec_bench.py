"""sketch_vector timing: Gaussian DenseSkOp<float> 1024 x 65536 applied to one vector, GEMV kernels vs the SIMT kernel."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, randblas_b200 as rb
d, m = 1024, 65536
x = torch.randn(m, device="cuda"); y = torch.zeros(d, device="cuda")
S = rb.DenseSkOp(rb.DenseDist(d, m), 0)
def t():
    for _ in range(2): rb.sketch_vector(rb.Op.NoTrans, 1.0, S, x, 1, 0.0, y, 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); rb.sketch_vector(rb.Op.NoTrans, 1.0, S, x, 1, 0.0, y, 1); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)
a = t(); ya = y.clone()
os.environ["RBX_NO_GEMV"] = "1"; b = t()
print(json.dumps({"gemv_ms": a, "simt_ms": b, "gsamples_per_s": d * m / a / 1e6, "max_abs_diff": float((ya - y).abs().max())}))
