#!/usr/bin/env python
"""Small shapes of every kernel family for compute-sanitizer (memcheck / racecheck on small shapes):
  compute-sanitizer --tool memcheck  --log-file gpurun_out/sanitizer_memcheck.log  python tools/sanitize_small.py
  compute-sanitizer --tool racecheck --log-file gpurun_out/sanitizer_racecheck.log python tools/sanitize_small.py
The SASO tile kernel (producer warp group, mbarriers, setmaxnreg), the tcgen05 sketch (TMA, TMEM, st.async into the peer CTA),
the DMMA sketch, the GEMV kernels, the sparse-data gather and the fills all run once; results are printed as checksums."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import randblas_b200 as rb  # noqa: E402

dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
only = set(sys.argv[1:])


def want(name):
    return not only or name in only


if want("fill"):
    buf = torch.empty(64 * 1001, dtype=torch.float32, device=dev)
    rb.fill_dense(rb.DenseDist(64, 1001), buf, rb.RNGState(1))
    print("fill_dense", float(buf.double().sum()))
if want("saso"):
    for dt in (torch.float64, torch.float32):
        for lay in ("C", "R"):
            d, n, m = 2048, 40, 2048 + 64
            A = torch.randn(m * n, dtype=dt, device=dev)
            B = torch.zeros(d * n, dtype=dt, device=dev)
            S = rb.SparseSkOp(rb.SparseDist(d, m, 8), 1997, dtype=dt)
            lda, ldb = (m, d) if lay == "C" else (n, n)
            rb.sketch_general(L(lay), O.NoTrans, O.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
            print("saso", dt, lay, float(B.double().abs().sum()))
    # LASO / general path
    A = torch.randn(300 * 20, dtype=torch.float64, device=dev)
    B = torch.zeros(16 * 20, dtype=torch.float64, device=dev)
    rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, 16, 20, 300, 1.0, rb.SparseSkOp(rb.SparseDist(16, 300, 4, "L"), 3), 0, 0, A, 300, 0.0, B, 16)
    print("laso", float(B.abs().sum()))
if want("tc"):
    d, n, m = 128, 256, 4096
    for fam in ("G", "U"):
        A = torch.randn(m * n, dtype=torch.float32, device=dev)
        B = torch.zeros(d * n, dtype=torch.float32, device=dev)
        rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, rb.DenseSkOp(rb.DenseDist(d, m, fam), 0), 0, 0, A, m, 0.0, B, d)
        print("tcgen05", fam, float(B.double().abs().sum()))
if want("dmma"):
    d, n, m = 128, 256, 4096
    A = torch.randn(m * n, dtype=torch.float64, device=dev)
    B = torch.zeros(d * n, dtype=torch.float64, device=dev)
    rb.sketch_general(L.ColMajor, O.NoTrans, O.NoTrans, d, n, m, 1.0, rb.DenseSkOp(rb.DenseDist(d, m), 0, dtype=torch.float64), 0, 0, A, m, 0.0, B, d)
    print("dmma", float(B.abs().sum()))
if want("gemv"):
    x = torch.randn(5000, dtype=torch.float64, device=dev)
    y = torch.zeros(300, dtype=torch.float64, device=dev)
    rb.sketch_vector(O.NoTrans, 300, 5000, 1.0, rb.DenseSkOp(rb.DenseDist(300, 5000), 5, dtype=torch.float64), 0, 0, x, 1, 0.0, y, 1)
    print("gemv", float(y.abs().sum()))
if want("spdata"):
    mm, nn, per, d = 2000, 1500, 6, 64
    g = torch.Generator().manual_seed(3)
    colidx = torch.sort(torch.randint(0, nn, (mm, per), generator=g), dim=1).values.reshape(-1).to(torch.int64)
    rowptr = torch.arange(0, mm * per + 1, per, dtype=torch.int64)
    vals = torch.randn(mm * per, generator=g, dtype=torch.float64)
    Asp = rb.CSRMatrix(mm, nn, vals.to(dev), rowptr.to(dev), colidx.to(dev))
    Bs = torch.zeros(d * nn, dtype=torch.float64, device=dev)
    rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, nn, mm, 1.0, rb.DenseSkOp(rb.DenseDist(d, mm), 9, dtype=torch.float64), 0, 0, Asp, 0, 0, 0.0, Bs, d)
    print("spdata", float(Bs.abs().sum()))
torch.cuda.synchronize()
print("done")
