"""fill_sparse over its parameter space: short / long major axis, vec_nnz from 1 to the full short axis, wide and tall, both index
types: time and nonzeros per second"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
dev = torch.device("cuda", 0)
for (r, c, k, ma, it) in [(2048, 1 << 20, 8, "S", torch.int64), (2048, 1 << 20, 8, "S", torch.int32), (1 << 20, 2048, 8, "S", torch.int64),
                          (2048, 1 << 20, 1, "S", torch.int64), (2048, 1 << 20, 32, "S", torch.int64), (2048, 1 << 18, 128, "S", torch.int64),
                          (2048, 1 << 16, 2048, "S", torch.int64), (65536, 1 << 16, 8, "S", torch.int64), (100, 1 << 20, 8, "S", torch.int64),
                          (2048, 1 << 20, 8, "L", torch.int64), (2048, 1 << 20, 4096, "L", torch.int64), (1 << 20, 2048, 4096, "L", torch.int64),
                          (64, 1 << 20, 1 << 17, "L", torch.int64)]:
    S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), 3, dtype=torch.float64, index_dtype=it)
    l0 = rb.lib.rbx_launch_count(); rb.fill_sparse(S); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); rb.fill_sparse(S); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[1]
    print(f"{r:8d} x {c:8d} vec_nnz {k:7d} {ma} {str(it)[6:]:6s} nnz {rb.nnz(S):10d} launches {nl} {t:8.3f} ms {rb.nnz(S) / t / 1e6:8.2f} Gnnz/s", flush=True)
