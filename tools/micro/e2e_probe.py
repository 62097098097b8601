"""Time the host-pointer SASO sketch (bench.py's e2e leg) per layout, with and without the slab pipeline."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, randblas_b200 as rb
D, n, m = 2048, 2048, 1 << 20
Ah = torch.empty(m * n, dtype=torch.float64, pin_memory=True); Ah.normal_()
Bh = torch.zeros(D * n, dtype=torch.float64, pin_memory=True)
S = rb.SparseSkOp(rb.SparseDist(D, m, 8), 1997)
for lay, lda, ldb in (("C", m, D), ("R", n, n)):
    for noslab in ("", "1"):
        if noslab: os.environ["RBX_NO_SLABS"] = "1"
        else: os.environ.pop("RBX_NO_SLABS", None)
        ts = []
        for i in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            rb.sketch_general(rb.Layout(lay), rb.Op.NoTrans, rb.Op.NoTrans, D, n, m, 1.0, S, 0, 0, Ah, lda, 0.0, Bh, ldb)
            ts.append(time.perf_counter() - t0)
        print(json.dumps({"layout": lay, "no_slabs": bool(noslab), "ms": [round(t * 1e3, 1) for t in ts]}), flush=True)
