import os, sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
os.environ["RBX_DENSE_MIN_MACS"] = "0"
import numpy as np, torch
import randblas_b200 as rb
import test_gpu_fuzz as F
from test_gpu_sketch import run_dense_case
cuda = torch.device("cuda:0")
for block in (26, 41):
    rng = np.random.default_rng(1000 + block)
    for it in range(14):
        pick, logint = F.pick, F.logint
        dt, side, layout = pick(rng, ["f32", "f64"]), pick(rng, "LR"), pick(rng, "CR")
        opS, opA = pick(rng, "NNT"), pick(rng, "NNT")
        d, n, m = logint(rng, 1, 300), logint(rng, 1, 300), logint(rng, 1 if rng.random() < 0.4 else 256, 6000)
        if rng.random() < 0.2:
            d, m = m % 700 + 1, d
        if side == "R":
            n, m = m, n
        alpha, beta = pick(rng, F.SCALARS)
        draw = dict(dt=dt, side=side, layout=layout, opS=opS, opA=opA, d=d, n=n, m=m, alpha=alpha, beta=beta, off=F.small_offsets(rng),
                    D_extra=F.small_offsets(rng), fam=pick(rng, "GU"), ma=pick(rng, "LS"), lda_pad=pick(rng, [0, 0, 1, 2, 3, 4, 8]),
                    ldb_pad=pick(rng, [0, 0, 1, 3, 4]), key=(int(rng.integers(0, 2 ** 32)), int(rng.integers(0, 2 ** 32))),
                    prefilled=bool(rng.random() < 0.2))
        l0 = rb.lib.rbx_launch_count()
        run_dense_case(rb, cuda, **draw)
        print(block, it, rb.lib.rbx_launch_count() - l0, {k: v for k, v in draw.items() if k != "key"}, flush=True)
