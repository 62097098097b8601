"""sparse-operator sketches off the streaming kernel's main road: lifts (B = S^T A with a wide S, B = S A with a tall S), transposed
data, right sketches in both layouts. d = 2048, m = 2^18, n = 512, double, vec_nnz = 8: launches, time, and the time per GB of
the larger dense operand as a yardstick."""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
d, m, n = 2048, 1 << 18, 512
dev = torch.device("cuda", 0)
L, O = rb.Layout, rb.Op
big = torch.empty(m * n, dtype=torch.float64, device=dev); rb.fill_dense(rb.DenseDist(m, n), big, rb.RNGState(1))
small = torch.empty(d * n, dtype=torch.float64, device=dev); rb.fill_dense(rb.DenseDist(d, n), small, rb.RNGState(2))
Bs = torch.zeros(d * n, dtype=torch.float64, device=dev)
Bb = torch.zeros(m * n, dtype=torch.float64, device=dev)
wide = rb.SparseSkOp(rb.SparseDist(d, m, 8), 7)
tall = rb.SparseSkOp(rb.SparseDist(m, d, 8), 7)
cases = []
for lay in (L.ColMajor, L.RowMajor):
    c = lay == L.ColMajor
    cases += [
        (f"sketch  S A          {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.NoTrans, O.NoTrans, d, n, m, 1.0, wide, 0, 0, big, m if c else n, 0.0, Bs, d if c else n)),
        (f"sketch  S A^T        {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.NoTrans, O.Trans, d, n, m, 1.0, wide, 0, 0, big, n if c else m, 0.0, Bs, d if c else n)),
        (f"sketch  tall^T A     {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.Trans, O.NoTrans, d, n, m, 1.0, tall, 0, 0, big, m if c else n, 0.0, Bs, d if c else n)),
        (f"right   A^T tall     {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.Trans, O.NoTrans, n, d, m, 1.0, big, m if c else n, tall, 0, 0, 0.0, Bs, n if c else d)),
        (f"right   A' tall      {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.NoTrans, O.NoTrans, n, d, m, 1.0, big, n if c else m, tall, 0, 0, 0.0, Bs, n if c else d)),
        (f"lift    wide^T A     {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.Trans, O.NoTrans, m, n, d, 1.0, wide, 0, 0, small, d if c else n, 0.0, Bb, m if c else n)),
        (f"lift    tall A       {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.NoTrans, O.NoTrans, m, n, d, 1.0, tall, 0, 0, small, d if c else n, 0.0, Bb, m if c else n)),
        (f"lift    A wide (R)   {lay.name}", lambda lay=lay, c=c: rb.sketch_general(lay, O.NoTrans, O.NoTrans, n, m, d, 1.0, small, n if c else d, wide, 0, 0, 0.0, Bb, n if c else m)),
    ]
for name, f in cases:
    l0 = rb.lib.rbx_launch_count(); f(); torch.cuda.synchronize(); nl = rb.lib.rbx_launch_count() - l0; ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); e0.record(); f(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"{name:34s} launches {nl:2d} {sorted(ts)[1]:8.3f} ms", flush=True)
