"""one fill_sparse shape for ncu: python tools/micro/fy_one.py ROWS COLS VEC_NNZ MAJOR_AXIS"""
import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import randblas_b200 as rb
r, c, k, ma = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), 3, dtype=torch.float64)
for _ in range(3):
    rb.fill_sparse(S)
torch.cuda.synchronize()
