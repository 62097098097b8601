"""sketch_vector around its kernels (d = 2048, m = 2^20): dense / sparse operators, NoTrans / Trans, both precisions, strided x"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
d, m = 2048, 1 << 20
dev = torch.device("cuda", 0)
for tdt in (torch.float32, torch.float64):
    x = torch.randn(2 * m, dtype=tdt, device=dev); y = torch.zeros(2 * m, dtype=tdt, device=dev)
    Sd = rb.DenseSkOp(rb.DenseDist(d, m), 3, dtype=tdt)
    Ss = rb.SparseSkOp(rb.SparseDist(d, m, 8), 3, dtype=tdt)
    for name, f in (("dense N", lambda: rb.sketch_vector(rb.Op.NoTrans, d, m, 1.0, Sd, 0, 0, x, 1, 0.0, y, 1)),
                    ("dense N incx 2", lambda: rb.sketch_vector(rb.Op.NoTrans, d, m, 1.0, Sd, 0, 0, x, 2, 0.0, y, 1)),
                    ("dense T", lambda: rb.sketch_vector(rb.Op.Trans, d, m, 1.0, Sd, 0, 0, x, 1, 0.0, y, 1)),
                    ("sparse N", lambda: rb.sketch_vector(rb.Op.NoTrans, d, m, 1.0, Ss, 0, 0, x, 1, 0.0, y, 1)),
                    ("sparse T", lambda: rb.sketch_vector(rb.Op.Trans, d, m, 1.0, Ss, 0, 0, x, 1, 0.0, y, 1))):
        l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        print(f"{str(tdt)[6:]:8s} {name:16s} launches {nl:2d} {sorted(ts)[1]:8.3f} ms", flush=True)
