"""dense sketches whose operands the 16-byte copies of the tensor-core kernels cannot read directly: double data with an odd
leading dimension, a pre-filled operator with an odd row length, a pre-filled short-axis-major operator (unit stride along
the sketch dimension). Prints launches and the median time per case; RBX_NO_TC=1 in the environment times the SIMT
fallback these cases used to take."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb

dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op


def timed(f):
    l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return nl, sorted(ts)[2]


for tdt, d, n in ((torch.float64, 1024, 512), (torch.float32, 2048, 2048)):
    name = str(tdt)[6:]
    for m, pad in ((65536, 0), (65537, 0), (65536, 1)):
        lda = m + pad
        A = torch.empty(lda * n, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(lda, n), A, rb.RNGState(99))
        C = torch.zeros(d * n, dtype=tdt, device=dev)
        S = rb.DenseSkOp(rb.DenseDist(d, m), 0, dtype=tdt)
        nl, t = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, C, d))
        print(name, "implicit  ColMajor m", m, "lda", lda, "launches", nl, round(t, 3), "ms", flush=True)
        # RowMajor data with an odd row length
        nn = n + (1 if (m % 2 or pad) else 0)
        Ar = torch.empty(m * nn, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(m, nn), Ar, rb.RNGState(98))
        Cr = torch.zeros(d * nn, dtype=tdt, device=dev)
        nl, t = timed(lambda: rb.sketch_general(L.RowMajor, O.NoTrans, O.NoTrans, d, nn, m, 1.0, S, 0, 0, Ar, nn, 0.0, Cr, nn))
        print(name, "implicit  RowMajor m", m, "n = lda", nn, "launches", nl, round(t, 3), "ms", flush=True)
        for ma in "LS":
            Sp = rb.DenseSkOp(rb.DenseDist(d, m, "G", ma), 0, dtype=tdt); rb.fill_dense(Sp)
            nl, t = timed(lambda: rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, Sp, 0, 0, A, lda, 0.0, C, d))
            print(name, "prefilled major axis", ma, "m", m, "lda", lda, "launches", nl, round(t, 3), "ms", flush=True)
