"""one sparse-operator sketch on the general path for a launch list: python tools/micro/general_path_one.py VEC_NNZ MAJOR_AXIS [LAYOUT]"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
k, ma = int(sys.argv[1]), sys.argv[2]
lay = sys.argv[3] if len(sys.argv) > 3 else "C"
d, m, n = 2048, 1 << 18, 2048
dev = torch.device("cuda", 0)
A = torch.empty(m * n, dtype=torch.float64, device=dev); rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(1))
B = torch.zeros(d * n, dtype=torch.float64, device=dev)
S = rb.SparseSkOp(rb.SparseDist(d, m, k, ma), 7)
for _ in range(2):
    rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m if lay == "C" else n, 0.0, B, d if lay == "C" else n)
torch.cuda.synchronize()
