"""sketch_sparse around C4 at 1/5 size (A 2e5 x 2e5, 100 nnz per row, d = 512): formats, sides, transposes, precisions"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
n, d, per_row = 200000, 512, 100
dev = torch.device("cuda", 0)
nnz = n * per_row
cols = torch.empty(nnz, dtype=torch.int64, device=dev)
rb.repeated_fisher_yates(rb.RNGState(7), per_row, n, n, cols)
cols = cols.view(n, per_row).sort(dim=1).values.reshape(-1).contiguous()
rowptr = torch.arange(0, nnz + 1, per_row, dtype=torch.int64, device=dev)
L, O = rb.Layout, rb.Op
for tdt in (torch.float64, torch.float32):
    vals = torch.empty(nnz, dtype=tdt, device=dev); rb.fill_dense(rb.DenseDist(nnz, 1), vals, rb.RNGState(8))
    csr = rb.CSRMatrix(n, n, vals, rowptr, cols)
    csc = rb.csr_to_csc(csr)
    rows = torch.arange(n, device=dev).repeat_interleave(per_row)
    coo = rb.COOMatrix(n, n, vals, rows, cols)
    B = torch.zeros(d * n, dtype=tdt, device=dev)
    S = rb.DenseSkOp(rb.DenseDist(d, n), 0, dtype=tdt)
    St = rb.DenseSkOp(rb.DenseDist(n, d), 0, dtype=tdt)
    cases = [("L CSR N", lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, csr, 0, 0, 0.0, B, d)),
             ("L CSC N", lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, csc, 0, 0, 0.0, B, d)),
             ("L COO N", lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, coo, 0, 0, 0.0, B, d)),
             ("L CSR T", lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.Trans, d, n, n, 1.0, S, 0, 0, csr, 0, 0, 0.0, B, d)),
             ("L CSR N RowMajor", lambda: rb.sketch_sparse(L.RowMajor, O.NoTrans, O.NoTrans, d, n, n, 1.0, S, 0, 0, csr, 0, 0, 0.0, B, n)),
             ("R CSR N", lambda: rb.sketch_sparse(L.ColMajor, O.NoTrans, O.NoTrans, n, d, n, 1.0, csr, 0, 0, St, 0, 0, 0.0, B, n)),
             ("R CSC N RowMajor", lambda: rb.sketch_sparse(L.RowMajor, O.NoTrans, O.NoTrans, n, d, n, 1.0, csc, 0, 0, St, 0, 0, 0.0, B, d))]
    for name, f in cases:
        l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        print(f"{str(tdt)[6:]:8s} {name:18s} launches {nl:2d} {sorted(ts)[1]:8.3f} ms", flush=True)
