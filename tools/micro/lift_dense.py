"""dense "lifts": a tall operator S (M x K, K small) applied to K x n data, B = S A -- the contraction axis is the SHORT one.
Times the public call for K = 16 .. 512 in both precisions and prints the launches (1 = SIMT kernel)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb

dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
M, n = 1 << 18, 1024
for tdt in (torch.float32, torch.float64):
    for K in (16, 32, 64, 128, 200, 256, 512):
        for layout in (L.ColMajor, L.RowMajor):
            A = torch.empty(K * n, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(K, n), A, rb.RNGState(99))
            B = torch.zeros(M * n, dtype=tdt, device=dev)
            S = rb.DenseSkOp(rb.DenseDist(M, K), 0, dtype=tdt)
            lda, ldb = (K, M) if layout == L.ColMajor else (n, n)
            f = lambda: rb.sketch_general(layout, O.NoTrans, O.NoTrans, M, n, K, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
            l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
            t = sorted(ts)[2]
            print(str(tdt)[6:], "K", K, str(layout).split(".")[-1], "launches", nl, round(t, 3), "ms", round(2.0 * M * n * K / t / 1e9, 1), "TFLOP/s", flush=True)
