"""The end-to-end callers of the hot path (reference examples/total-least-squares and
examples/sparse-low-rank-approx) run on the GPU at small sizes and their own acceptance quantities are checked."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_example(name, *args):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", name), *map(str, args)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


@pytest.mark.gpu
def test_tls_sketch_and_solve_example():
    """tls_sparse_skop.cc / tls_dense_skop.cc: the sketched solution stays close to the classical one"""
    for extra in ((), ("--dense",)):
        r = run_example("tls_sketch_and_solve.py", 40000, 40, *extra)
        assert 0 <= r["rel_distance_to_classical"] < 0.5 and r["x_minus_ones_rel"] < 0.5, r


@pytest.mark.gpu
def test_svd_sparse_rank1_plus_noise_example():
    """svd_rank1_plus_noise.cc: the planted singular vectors are recovered"""
    r = run_example("svd_sparse_rank1_plus_noise.py", 20000, 10000, 4)
    assert r["rel_angle_left"] < 1e-4 and r["rel_angle_right"] < 1e-4, r


@pytest.mark.gpu
def test_qrcp_sparse_example(tmp_path):
    """qrcp_matrixmarket.cc: Q orthonormal, A(:, piv[:k]) = Q R(:, :k) with R upper triangular -- on a synthetic matrix and on
    a MatrixMarket file"""
    r = run_example("qrcp_sparse.py", 12, "--m", 20000, "--n", 6000)
    assert r["orthogonality_error"] < 1e-8 and r["rel_residual_pivot_cols"] < 1e-8 and r["upper_triangular_defect"] < 1e-8, r
    assert r["rel_residual_first_4k_cols"] < 0.9, r
    import numpy as np
    rng = np.random.default_rng(0)
    m, n, nnz = 300, 120, 2500
    rows, cols = rng.integers(1, m + 1, nnz), rng.integers(1, n + 1, nnz)
    _, first = np.unique(rows * (n + 1) + cols, return_index=True)
    path = tmp_path / "a.mtx"
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n% synthetic\n")
        f.write(f"{m} {n} {len(first)}\n")
        for i in first:
            f.write(f"{rows[i]} {cols[i]} {rng.standard_normal():.17g}\n")
    r = run_example("qrcp_sparse.py", 8, path)
    assert r["m"] == m and r["n"] == n and r["nnz"] == len(first), r
    assert r["orthogonality_error"] < 1e-8 and r["rel_residual_pivot_cols"] < 1e-8, r
