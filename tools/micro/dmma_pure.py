"""C5 slab shape (d = 8192, m = 65536, n = 4096, double): implicit operator (panel fills overlapped on the auxiliary stream) against a
pre-filled operator (pure DMMA GEMM, one launch) and cuBLAS DGEMM on the same shape"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
d, m, n = 8192, 65536, 4096
A = torch.empty(m * n, dtype=torch.float64, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
B = torch.zeros(d * n, dtype=torch.float64, device=dev)
def timed(f, reps=3):
    f(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
S = rb.DenseSkOp(rb.DenseDist(d, m), 0, dtype=torch.float64)
t = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d))
print("implicit ", round(t, 2), "ms", round(2.0 * d * m * n / t / 1e9, 2), "TFLOP/s", flush=True)
os.environ["RBX_DMMA_NO_OVERLAP"] = "1"
t = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d))
print("implicit, fills not overlapped", round(t, 2), "ms", round(2.0 * d * m * n / t / 1e9, 2), "TFLOP/s", flush=True)
del os.environ["RBX_DMMA_NO_OVERLAP"]
rb.fill_dense(S)
t = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d))
print("prefilled", round(t, 2), "ms", round(2.0 * d * m * n / t / 1e9, 2), "TFLOP/s", flush=True)
Sm = S.buff.view(d, m) if S.buff.numel() == d * m else None
Am = A.view(n, m).t()
t = timed(lambda: torch.matmul(Sm, Am))
print("cuBLAS   ", round(t, 2), "ms", round(2.0 * d * m * n / t / 1e9, 2), "TFLOP/s", flush=True)
