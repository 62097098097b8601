"""fill_sparse, vec_nnz = 9 .. 64 on a 2048 x 2^20 SASO: the per-thread lazy-list kernel against the per-CTA hash kernel
(RBX_FY_HASH_MIN selects where the second takes over)"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
for hm in ("1000000", "0"):
    os.environ["RBX_FY_HASH_MIN"] = hm
    for k in (4, 8, 12, 16, 24, 32, 48, 64):
        S = rb.SparseSkOp(rb.SparseDist(2048, 1 << 20, k), 3, dtype=torch.float64)
        rb.fill_sparse(S); torch.cuda.synchronize(); ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); rb.fill_sparse(S); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        print("hash" if hm == "0" else "lazy", "vec_nnz", k, round(sorted(ts)[1], 3), "ms", round(rb.nnz(S) / sorted(ts)[1] / 1e6, 1), "Gnnz/s", flush=True)
