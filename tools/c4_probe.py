import os, sys, json, time
sys.path.insert(0, "/root/repo")
import torch, randblas_b200 as rb
dev = torch.device("cuda", 0)
n, d, per_row = 1000000, 512, 100
nnz = n * per_row
cols = torch.empty(nnz, dtype=torch.int64, device=dev)
rb.repeated_fisher_yates(rb.RNGState(7), per_row, n, n, cols)
cols = cols.view(n, per_row).sort(dim=1).values.reshape(-1).contiguous()
vals = torch.empty(nnz, dtype=torch.float64, device=dev)
rb.fill_dense(rb.DenseDist(nnz, 1), vals, rb.RNGState(8))
rowptr = torch.arange(0, nnz + 1, per_row, dtype=torch.int64, device=dev)
A = rb.CSRMatrix(n, n, vals, rowptr, cols)
S = rb.DenseSkOp(rb.DenseDist(d, n), 0, dtype=torch.float64)
B = torch.zeros(d * n, dtype=torch.float64, device=dev)
def run():
    rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, n, 1.0, S, 0, 0, A, 0, 0, 0.0, B, d)
ref = None
DCS = sys.argv[1].split(",") if len(sys.argv) > 1 else ("0", "16", "8", "32")
REPS = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for dc in DCS:
    os.environ["RBX_SPD_DC"] = dc
    run(); torch.cuda.synchronize()
    ts = []
    for _ in range(REPS):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    chk = float(B.double().abs().sum())
    if ref is None: ref = B.clone()
    err = float((B - ref).abs().max() / ref.abs().max())
    print(json.dumps({"dc": dc, "ms": sorted(ts) if ts else None, "rel_diff_vs_dc0": err}), flush=True)
