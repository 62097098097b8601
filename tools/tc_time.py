import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import randblas_b200 as rb
d, n, m = [int(x) for x in os.environ.get("DNM", "1024,1024,65536").split(",")]
dev = torch.device("cuda", 0)
A = torch.empty(m * n, dtype=torch.float32, device=dev)
rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
C = torch.zeros(d * n, dtype=torch.float32, device=dev)
for fam in sys.argv[1:] or ["G", "U"]:
    S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 0)
    f = lambda: rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, C, d)
    f(); torch.cuda.synchronize()
    ts = []
    for _ in range(int(os.environ.get("REPS","8"))):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"dbg={os.environ.get('RBX_TC_DBG','0')} cl={os.environ.get('RBX_TC_CLUSTER','-')} ksplit={os.environ.get('RBX_TC_KSPLIT','-')} fam={fam}: median {t:.3f} ms  {2.0*d*n*m/t/1e9:.1f} TFLOP/s   all: " + " ".join(f"{x:.2f}" for x in ts), flush=True)
