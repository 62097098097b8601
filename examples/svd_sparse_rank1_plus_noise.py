#!/usr/bin/env python
"""Randomized SVD of a sparse matrix on the GPU: the reference's
examples/sparse-low-rank-approx/svd_rank1_plus_noise.cc:235-258 (qb_decompose_sparse_matrix + qb_to_svd) with every
RandBLAS call served by this library:

  A = signal_scale * u v^T (u, v with vec_nnz nonzeros each, positions from repeated_fisher_yates like make_signal_matrix)
      + sparse iid noise (density prob_nonzero), held as a COOMatrix on the device;
  S = fill_dense(DenseDist(n, k));  power iteration  S <- orth(A^T orth(A S))  (p = 2);  Q = orth(A S);  B = Q^T A;
  B = U diag(s) V^T (small SVD);  left vectors Q U.

left_spmm / right_spmm are this repository's output-stationary sparse x dense kernels (no atomics), fill_dense its bit-exact
generator; QR and the k x n SVD go through cuSOLVER (torch.linalg), as lapack::geqrf / gesdd do in the reference.
Prints one JSON line: run time and the relative angles between the recovered and the planted top singular vectors.

  python examples/svd_sparse_rank1_plus_noise.py [m n vec_nnz]
"""
import argparse
import json
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import randblas_b200 as rb  # noqa: E402

L, O = rb.Layout.ColMajor, rb.Op


def orth(M):
    return torch.linalg.qr(M, mode="reduced")[0]


def col(t, rows, cols):
    """(rows x cols) matrix view of a column-major flat tensor"""
    return t.view(cols, rows).t()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("m", nargs="?", type=int, default=200000)
    ap.add_argument("n", nargs="?", type=int, default=100000)
    ap.add_argument("vec_nnz", nargs="?", type=int, default=4)
    args = ap.parse_args()
    m, n, vn = args.m, args.n, args.vec_nnz
    dev = torch.device("cuda", 0)
    signal_scale, noise_scale, prob_nonzero = 1e2, 1e-6, 1e-4

    # signal: u, v with vec_nnz signed entries at positions drawn without replacement (make_signal_matrix)
    iu = torch.empty(vn, dtype=torch.int64, device=dev); vu = torch.empty(vn, dtype=torch.float64, device=dev)
    iv = torch.empty(vn, dtype=torch.int64, device=dev); vv = torch.empty(vn, dtype=torch.float64, device=dev)
    rb.repeated_fisher_yates(rb.RNGState(0), vn, m, 1, iu, None, vu)
    rb.repeated_fisher_yates(rb.RNGState(2), vn, n, 1, iv, None, vv)      # (the reference reuses the first state here: repeated_fisher_yates returns its input, sparse_skops.hh:105)
    u = torch.zeros(m, dtype=torch.float64, device=dev); v = torch.zeros(n, dtype=torch.float64, device=dev)
    u[iu] = vu / math.sqrt(vn); v[iv] = vv / math.sqrt(vn)
    srows = iu.repeat(vn); scols = iv.repeat_interleave(vn)
    svals = signal_scale * (u[srows] * v[scols])
    # noise: iid positions at the given density, uniform values (iid_sparsify_random_dense), generated on the device
    nnz_noise = int(m * n * prob_nonzero)
    g = torch.Generator(device=dev).manual_seed(1)
    nrows = torch.randint(0, m, (nnz_noise,), generator=g, device=dev)
    ncols = torch.randint(0, n, (nnz_noise,), generator=g, device=dev)
    nvals = torch.empty(nnz_noise, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(nnz_noise, 1, "U"), nvals, rb.RNGState(1))
    A = rb.COOMatrix(m, n, torch.cat([svals, noise_scale * nvals]), torch.cat([srows, nrows]), torch.cat([scols, ncols]))

    k, p = max(3, vn), 2
    Q = torch.empty(m * k, dtype=torch.float64, device=dev)
    S = torch.empty(n * k, dtype=torch.float64, device=dev)
    B = torch.empty(k * n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rb.fill_dense(rb.DenseDist(n, k), S, rb.RNGState(0))                       # natural layout of a tall DenseDist is ColMajor
    for _ in range(p // 2):                                                   # S = orth(A' * orth(A * S))
        rb.left_spmm(L, O.NoTrans, O.NoTrans, m, k, n, 1.0, A, 0, 0, S, n, 0.0, Q, m)
        col(Q, m, k).copy_(orth(col(Q, m, k)))
        rb.left_spmm(L, O.Trans, O.NoTrans, n, k, m, 1.0, A, 0, 0, Q, m, 0.0, S, n)
        col(S, n, k).copy_(orth(col(S, n, k)))
    rb.left_spmm(L, O.NoTrans, O.NoTrans, m, k, n, 1.0, A, 0, 0, S, n, 0.0, Q, m)
    col(Q, m, k).copy_(orth(col(Q, m, k)))
    rb.right_spmm(L, O.Trans, O.NoTrans, k, n, m, 1.0, Q, m, A, 0, 0, 0.0, B, k)   # B = Q' A
    U, s, Vh = torch.linalg.svd(col(B, k, n), full_matrices=False)
    Utop = col(Q, m, k) @ U[:, 0]
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    cu, cv = float(torch.dot(u, Utop).abs().clamp(max=1.0)), float(torch.dot(v, Vh[0]).abs().clamp(max=1.0))
    print(json.dumps({"m": m, "n": n, "vec_nnz": vn, "nnz": A.nnz, "k": k, "power_iterations": p, "runtime_s": dt,
                      "sigma_top": float(s[0]), "rel_angle_left": math.acos(cu) / math.pi, "rel_angle_right": math.acos(cv) / math.pi}), flush=True)


if __name__ == "__main__":
    main()
