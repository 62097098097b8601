#!/usr/bin/env python
"""Truncated column-pivoted QR of a sparse matrix on the GPU: the reference's
examples/sparse-low-rank-approx/qrcp_matrixmarket.cc:157-291 (power_iter_col_sketch + sketch_to_tqrcp) with every RandBLAS
call served by this library:

  Y (k x n) = S A, S built with p passes over A:  S0 Gaussian from fill_dense, then alternately  T = row_orth(S A),
      S = row_orth(T A^T)  (right_spmm with the CSC matrix, rows stabilised by a Householder LQ like qr_row_stabilize);
  piv = pivots of GEQP3(Y);  Q = A(:, piv[:k]) gathered densely, preconditioned with R = triu(Y_qr)[:k, :k] and
      Cholesky-orthogonalised;  R = (Q^T A)(:, piv)  (right_spmm again).

right_spmm / fill_dense are this repository's kernels (output-stationary gather over device-regrouped columns, bit-exact
generator). The k x n GEQP3 runs on the host through LAPACK (scipy.linalg.qr(pivoting=True)) -- cuSOLVER has no pivoted QR
and the reference calls LAPACK for it too; the tall QR / Cholesky / triangular solves go through cuSOLVER / cuBLAS (torch).
Prints one JSON line: run time, the leading pivots and ||A(:, piv) - Q R||_F / ||A||_F over the first 4k pivoted columns.

  python examples/qrcp_sparse.py [k] [matrix.mtx]        (no file: a synthetic m x n matrix of rank ~ k plus sparse noise)
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import randblas_b200 as rb  # noqa: E402

L, O = rb.Layout.ColMajor, rb.Op


def col(t, rows, cols):
    """(rows x cols) matrix view of a column-major flat tensor"""
    return t.view(cols, rows).t()


def read_matrix_market(path, dev):
    """coordinate real general / symmetric files, the subset qrcp_matrixmarket.cc's fast_matrix_market call is used for"""
    with open(path) as f:
        head = f.readline().lower().split()
        if head[:3] != ["%%matrixmarket", "matrix", "coordinate"]:
            raise ValueError("only coordinate MatrixMarket files are supported")
        sym = head[4] == "symmetric"
        line = f.readline()
        while line.startswith("%"):
            line = f.readline()
        m, n, _ = (int(x) for x in line.split())
        data = np.loadtxt(f, ndmin=2)
    r, c = data[:, 0].astype(np.int64) - 1, data[:, 1].astype(np.int64) - 1
    v = data[:, 2] if data.shape[1] > 2 else np.ones(len(r))
    if sym:
        off = r != c
        r, c, v = np.concatenate([r, c[off]]), np.concatenate([c, r[off]]), np.concatenate([v, v[off]])
    t = lambda a, dt: torch.as_tensor(a, dtype=dt, device=dev)
    return rb.COOMatrix(m, n, t(v, torch.float64), t(r, torch.int64), t(c, torch.int64))


def synthetic(m, n, k, dev):
    """k sparse rank-one terms with decaying weights plus sparse noise: the pivots should pick the heavy columns"""
    g = torch.Generator(device=dev).manual_seed(7)
    rows, cols, vals = [], [], []
    for t in range(k):
        ru = torch.randint(0, m, (40,), generator=g, device=dev); cu = torch.randint(0, n, (40,), generator=g, device=dev)
        rows.append(ru.repeat(40)); cols.append(cu.repeat_interleave(40))
        vals.append(torch.full((1600,), 10.0 * 0.8 ** t, dtype=torch.float64, device=dev))
    nnz_noise = int(1e-4 * m * n)
    rows.append(torch.randint(0, m, (nnz_noise,), generator=g, device=dev)); cols.append(torch.randint(0, n, (nnz_noise,), generator=g, device=dev))
    nv = torch.empty(nnz_noise, dtype=torch.float64, device=dev)
    rb.fill_dense(rb.DenseDist(nnz_noise, 1, "U"), nv, rb.RNGState(1))
    vals.append(1e-3 * nv)
    return rb.COOMatrix(m, n, torch.cat(vals), torch.cat(rows), torch.cat(cols))


def row_orth(M):
    """qr_row_stabilize (qrcp_matrixmarket.cc:101-114): replace the k x n matrix by the orthonormal row basis of its LQ"""
    Qt = torch.linalg.qr(M.t(), mode="reduced")[0]
    M.copy_(Qt.t())


def power_iter_col_sketch(A, k, p, state, dev):
    """qrcp_matrixmarket.cc:213-297: Y = S A with p data-aware passes"""
    m, n = A.n_rows, A.n_cols
    W1 = torch.empty(k * n, dtype=torch.float64, device=dev)      # mat_work1 = Y (k x n)
    W2 = torch.empty(k * m, dtype=torch.float64, device=dev)      # mat_work2 (k x m)
    done = 0
    if p % 2 == 0:
        rb.fill_dense(rb.DenseDist(k, m), W2, state)              # (read as k x m ColMajor below, like the reference: iid entries either way)
    else:
        rb.fill_dense(rb.DenseDist(k, n), W1, state)
        rb.right_spmm(L, O.NoTrans, O.Trans, k, m, n, 1.0, W1, k, A, 0, 0, 0.0, W2, k)
        done += 1
        row_orth(col(W2, k, m))
    while p - done > 0:
        rb.right_spmm(L, O.NoTrans, O.NoTrans, k, n, m, 1.0, W2, k, A, 0, 0, 0.0, W1, k)
        done += 1
        row_orth(col(W1, k, n))
        rb.right_spmm(L, O.NoTrans, O.Trans, k, m, n, 1.0, W1, k, A, 0, 0, 0.0, W2, k)
        done += 1
        row_orth(col(W2, k, m))
    rb.right_spmm(L, O.NoTrans, O.NoTrans, k, n, m, 1.0, W2, k, A, 0, 0, 0.0, W1, k)
    return W1


def gather_columns(A, cols, out):
    """out (m x len(cols), zero on entry) = A(:, cols) for a CSCMatrix A (duplicate entries add up, as they do in spmm)"""
    ptr = A.colptr.cpu()
    for j, cj in enumerate(cols.cpu().tolist()):
        lo, hi = int(ptr[cj]), int(ptr[cj + 1])
        out[:, j].index_put_((A.rowidxs[lo:hi],), A.vals[lo:hi], accumulate=True)


def sketch_to_tqrcp(A, k, Y, dev):
    """qrcp_matrixmarket.cc:321-385. Returns Q (m x k, orthonormal basis of A(:, piv[:k])), R = (Q^T A)(:, piv), piv."""
    import scipy.linalg
    m, n = A.n_rows, A.n_cols
    # step 1: pivots (GEQP3 on the host: k x n with small k)
    _, Rqr, piv = scipy.linalg.qr(col(Y, k, n).cpu().numpy(), mode="economic", pivoting=True)
    piv_d = torch.as_tensor(piv.astype(np.int64), device=dev)
    # step 2: A(:, piv[:k]) as a dense m x k matrix
    Q = torch.zeros(m, k, dtype=torch.float64, device=dev)
    gather_columns(A, piv_d[:k], Q)
    # step 3: orth(Q): precondition with the sketch's triangular factor, then Cholesky QR
    pre = torch.as_tensor(np.ascontiguousarray(np.triu(Rqr[:k, :k])), device=dev)
    Q = torch.linalg.solve_triangular(pre, Q, upper=True, left=False)
    G = Q.t() @ Q
    Q = torch.linalg.solve_triangular(torch.linalg.cholesky(G, upper=True), Q, upper=True, left=False)
    # step 4: R = Q^T A, columns pivoted
    Qf = Q.t().contiguous().view(-1)                              # m x k ColMajor flat
    rb.right_spmm(L, O.Trans, O.NoTrans, k, n, m, 1.0, Qf, m, A, 0, 0, 0.0, Y, k)
    R = col(Y, k, n)[:, piv_d]
    return Q, R, piv_d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("k", nargs="?", type=int, default=16)
    ap.add_argument("mtx", nargs="?", default=None)
    ap.add_argument("--m", type=int, default=200000)
    ap.add_argument("--n", type=int, default=50000)
    ap.add_argument("--power", type=int, default=2)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    coo = read_matrix_market(args.mtx, dev) if args.mtx else synthetic(args.m, args.n, args.k, dev)
    if coo.n_rows < coo.n_cols and args.mtx:
        coo = rb.COOMatrix(coo.n_cols, coo.n_rows, coo.vals, coo.cols, coo.rows)     # the reference transposes wide inputs
    A = rb.coo_to_csc(coo)                                                            # device-side sort + compression
    k = min(args.k, A.n_cols)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    Y = power_iter_col_sketch(A, k, args.power, rb.RNGState(0), dev)
    Q, R, piv = sketch_to_tqrcp(A, k, Y, dev)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    # check on the leading 4k pivoted columns: A(:, piv[:c]) ~ Q R(:, :c); the first k of them exactly
    c = min(4 * k, A.n_cols)
    Ap = torch.zeros(A.n_rows, c, dtype=torch.float64, device=dev)
    gather_columns(A, piv[:c], Ap)
    res = float(torch.linalg.norm(Ap - Q @ R[:, :c])) / float(torch.linalg.norm(A.vals))
    res_k = float(torch.linalg.norm(Ap[:, :k] - Q @ R[:, :k])) / max(float(torch.linalg.norm(Ap[:, :k])), 1e-300)
    orth_err = float(torch.linalg.norm(Q.t() @ Q - torch.eye(k, dtype=torch.float64, device=dev)))
    print(json.dumps({"m": A.n_rows, "n": A.n_cols, "nnz": A.nnz, "k": k, "power_iterations": args.power, "runtime_s": dt,
                      "leading_pivots": [int(x) for x in piv[: min(k, 8)]], "rel_residual_first_4k_cols": res,
                      "rel_residual_pivot_cols": res_k, "orthogonality_error": orth_err,
                      "upper_triangular_defect": float(torch.linalg.norm(torch.tril(R[:, :k], -1))) / float(torch.linalg.norm(R[:, :k]))}), flush=True)
    assert orth_err < 1e-8 and res_k < 1e-8 and math.isfinite(res)


if __name__ == "__main__":
    main()
