#!/usr/bin/env python
"""Sketch-and-solve total least squares on the GPU: the caller of the hot path in
examples/total-least-squares/tls_sparse_skop.cc:100-200 and tls_dense_skop.cc of the reference, end to end on one B200.

  [A | b] is m x (n + 1), m >> n. Classical TLS takes the SVD of the whole augmented matrix; the randomized method sketches
  it to 2 (n + 1) rows with a SASO (vec_nnz = 8, seed 1997, as the reference example) or a Gaussian operator and takes the SVD
  of the small matrix. x = -V[:n, n] / V[n, n]  (total_least_squares(), tls_sparse_skop.cc:70-91).

The sketch is this repository's sketch_general (hand-written kernels behind the C ABI); the small dense factorisation is
cuSOLVER's gesvd through torch.linalg.svd (a library call, as lapack::gesdd is in the reference). Data is the reference's
init_noisy_data: A Gaussian (key 0), b = A * 1 + Gaussian noise (key 1), generated on the device by fill_dense.

  python examples/tls_sketch_and_solve.py [m n] [--dense]
prints one JSON line with the times of each stage and the relative distance between the sketched and the classical solution.
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import randblas_b200 as rb  # noqa: E402


def tls_from_svd(M, n):
    """x from the right singular vector of the smallest singular value of the (rows x (n + 1)) matrix M"""
    _, s, Vh = torch.linalg.svd(M, full_matrices=False)
    v = Vh[n, :]
    return -v[:n] / v[n], s


def timed(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    return out, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("m", nargs="?", type=int, default=1000000)
    ap.add_argument("n", nargs="?", type=int, default=500)
    ap.add_argument("--dense", action="store_true", help="Gaussian DenseSkOp instead of the SASO")
    ap.add_argument("--skip-classical", action="store_true")
    args = ap.parse_args()
    m, n = args.m, args.n
    dev = torch.device("cuda", 0)
    sk_dim = 2 * (n + 1)
    # [A | b], column-major m x (n + 1): init_noisy_data (tls_sparse_skop.cc:50-68)
    AB = torch.empty(m * (n + 1), dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(m, n), AB[: m * n], rb.RNGState(0))
    eps = torch.empty(m, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(m, 1), eps, rb.RNGState(1))
    A = AB[: m * n].view(n, m).t()                      # m x n view of the column-major block
    AB[m * n:] = A.sum(dim=1) + eps                     # b = A * ones + noise
    ABm = AB.view(n + 1, m).t()                         # m x (n + 1)

    if args.dense:
        S = rb.DenseSkOp(rb.DenseDist(sk_dim, m), 1997, dtype=torch.float64)
    else:
        S = rb.SparseSkOp(rb.SparseDist(sk_dim, m, 8), 1997)
    SAB = torch.zeros(sk_dim * (n + 1), dtype=torch.float64, device=dev)

    def sketch():
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, sk_dim, n + 1, m, 1.0, S, 0, 0, AB, m, 0.0, SAB, sk_dim)

    sketch()                                            # warm-up (workspace pools, module load)
    _, t_sketch = timed(sketch)
    (x_sk, _), t_solve = timed(lambda: tls_from_svd(SAB.view(n + 1, sk_dim).t(), n))
    out = {"m": m, "n": n, "sk_dim": sk_dim, "operator": "Gaussian DenseSkOp<double>" if args.dense else "SASO SparseSkOp<double> vec_nnz=8",
           "sketch_s": t_sketch, "sketched_tls_s": t_solve, "randomized_total_s": t_sketch + t_solve,
           "sketch_gflops": (2.0 * (8 if not args.dense else sk_dim) * m * (n + 1)) / t_sketch / 1e9,
           "x_minus_ones_rel": float((x_sk - 1).norm() / (n ** 0.5))}
    if not args.skip_classical:
        tls_from_svd(ABm[: 4 * (n + 1)], n)             # warm up cuSOLVER
        (x_true, _), t_true = timed(lambda: tls_from_svd(ABm, n))
        out.update({"classical_tls_s": t_true, "speedup_vs_classical_gpu": t_true / (t_sketch + t_solve),
                    "rel_distance_to_classical": float((x_sk - x_true).norm() / x_true.norm())})
    print(json.dumps(out), flush=True)
    return out


if __name__ == "__main__":
    main()
