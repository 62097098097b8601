"""tcgen05 dense sketch vs the SIMT kernel (RBX_NO_TC=1) on the same inputs; prints max relative error and timings."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import randblas_b200 as rb

def run(d, n, m, fam, no_tc, reps=3):
    if no_tc: os.environ["RBX_NO_TC"] = "1"
    else: os.environ.pop("RBX_NO_TC", None)
    dev = torch.device("cuda", 0)
    A = torch.empty(m * n, dtype=torch.float32, device=dev)
    rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    B = torch.full((d * n,), 0.25, dtype=torch.float32, device=dev)
    S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 0)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, S, 0, 0, A, m, -0.5, B, d)
    torch.cuda.synchronize()
    C = torch.zeros(d * n, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, C, d)
    e1.record(); torch.cuda.synchronize()
    return B.double().cpu(), e0.elapsed_time(e1) / reps

for (d, n, m) in [(128, 256, 1024), (256, 512, 4096), (200, 300, 5000), (1024, 1024, 65536)]:
    for fam in ("G", "U"):
        b1, t1 = run(d, n, m, fam, False)
        b0, t0 = run(d, n, m, fam, True)
        err = float((b1 - b0).abs().max() / b0.abs().max())
        print(f"d={d} n={n} m={m} fam={fam}: relerr={err:.3e}  tc={t1:.3f} ms ({2.0*d*n*m/t1/1e9:.1f} TFLOP/s)  simt={t0:.3f} ms", flush=True)
