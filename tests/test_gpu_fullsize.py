"""The BASELINE.json configurations at their FULL sizes (the oracle cannot finish these in seconds), checked through
size-independent properties: exact rows / columns / entries of the result recomputed in FP64 from the operator that the
device fill produced (that fill is pinned bit for bit against the oracle and the reference in test_gpu_fill.py),
column-sum identities, linearity in (alpha, beta) and run-to-run bit-reproducibility.
C5 needs 268 GB for A: one 1/8 column slab (what one GPU of the 8-GPU configuration holds) is checked here; the sharded
forms are covered by tests/test_sharded_cpu.py and tools/bench_c5_sharded.py."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def need_gb(cuda, gb):
    free, _ = torch.cuda.mem_get_info(cuda)
    if free < gb * (1 << 30):
        pytest.skip(f"needs {gb} GB of device memory, {free >> 30} GB free")


@pytest.mark.parametrize("layout", ["C", "R"])
def test_c2_saso_full_size(rb, cuda, layout):
    """C2: SparseSkOp<double>, vec_nnz 8, d = 2048, A 2^20 x 2048 (lskges; test_lskges.cc at full size)."""
    need_gb(cuda, 24)
    d, n, m, k = 2048, 2048, 1 << 20, 8
    lda, ldb = (m, d) if layout == "C" else (n, n)
    A = torch.empty(m * n, dtype=torch.float64, device=cuda)
    rb.fill_dense(rb.Layout(layout), rb.DenseDist(m, n), m, n, 0, 0, A, rb.RNGState(99))
    Am = A.view(n, m).t() if layout == "C" else A.view(m, n)              # (m x n) view
    S = rb.SparseSkOp(rb.SparseDist(d, m, k), 1234)                       # not filled: the streaming kernel regenerates it
    Sf = rb.SparseSkOp(rb.SparseDist(d, m, k), 1234)
    rb.fill_sparse(Sf)
    B = torch.full((d * n,), float("nan"), dtype=torch.float64, device=cuda)
    L, O = rb.Layout(layout), rb.Op.NoTrans
    rb.sketch_general(L, O, O, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
    Bm = B.view(n, d).t() if layout == "C" else B.view(d, n)
    assert bool(torch.isfinite(B).all())
    # (1) column sums: 1^T (S A) = (1^T S) A, with 1^T S accumulated from the filled triplets
    s = torch.zeros(m, dtype=torch.float64, device=cuda).index_add_(0, Sf.cols, Sf.vals)
    want = s @ Am
    scale = float(k) * Am.abs().sum(dim=0)
    assert float(((Bm.sum(dim=0) - want).abs() / scale).max()) < 1e-12
    # (2) sixteen rows of B recomputed exactly from the triplets of those rows
    g = torch.Generator(device="cpu").manual_seed(5)
    for i in torch.randint(0, d, (16,), generator=g).tolist():
        sel = (Sf.rows == i).nonzero().squeeze(1)
        ref = (Sf.vals[sel].unsqueeze(1) * Am[Sf.cols[sel]]).sum(dim=0)
        mag = Am[Sf.cols[sel]].abs().sum(dim=0).clamp_min(1e-300)
        assert float(((Bm[i] - ref).abs() / mag).max()) < sel.numel() * 2.0 ** -52
    # (3) bit-reproducible, and B <- 2 S A + B is fl(3 B) exactly (2 x is exact; one rounding either way)
    B2 = torch.empty_like(B)
    rb.sketch_general(L, O, O, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B2, ldb)
    assert torch.equal(B, B2)
    rb.sketch_general(L, O, O, d, n, m, 2.0, S, 0, 0, A, lda, 1.0, B2, ldb)
    assert torch.equal(B2, 2.0 * B + B)


@pytest.mark.parametrize("fam", ["G", "U"])
def test_c1_dense_full_size(rb, cuda, fam):
    """C1: DenseSkOp<float> d = 1024, A 65536 x 1024 ColMajor (lskge3): against an FP64 product with the filled operator,
    componentwise bound of the reference's tests (linop_common.hh:257-263) with eps = 2^-23."""
    d, n, m = 1024, 1024, 65536
    A = torch.empty(m * n, dtype=torch.float32, device=cuda)
    rb.fill_dense(rb.DenseDist(m, n), A, rb.RNGState(99))
    S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 7)
    B = torch.full((d * n,), float("nan"), dtype=torch.float32, device=cuda)
    L, O = rb.Layout.ColMajor, rb.Op.NoTrans
    rb.sketch_general(L, O, O, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d)
    Sbuf = torch.empty(d * m, dtype=torch.float32, device=cuda)
    rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(d, m, fam), d, m, 0, 0, Sbuf, rb.RNGState(7))
    Sm = Sbuf.view(m, d).t().double()
    Am = A.view(n, m).t().double()
    exact = Sm @ Am
    bound = m * 2 * 2.0 ** -23 * (Sm.abs() @ Am.abs())
    ratio = float(((B.view(n, d).t().double() - exact).abs() / bound).max())
    assert ratio <= 1.0, ratio
    B2 = torch.empty_like(B)
    rb.sketch_general(L, O, O, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B2, d)
    assert torch.equal(B, B2)


def test_c3_fill_full_size(rb, cuda):
    """C3: fill_dense of 2^32 floats (65536 x 65536 Gaussian): windows regenerated with fill_dense's submatrix form must
    equal the same windows of the full buffer bit for bit (test_denseskop.cc:162-296), and the fill is reproducible."""
    need_gb(cuda, 36)
    n = 65536
    D = rb.DenseDist(n, n, "G", "L")
    full = torch.empty(n * n, dtype=torch.float32, device=cuda)
    rb.fill_dense(D, full, rb.RNGState(11))
    F = full.view(n, n) if rb.dist_to_layout(D) == rb.Layout.RowMajor else full.view(n, n).t()
    lay = rb.dist_to_layout(D)
    g = torch.Generator(device="cpu").manual_seed(3)
    for _ in range(12):
        r, c = (int(x) for x in torch.randint(1, 300, (2,), generator=g))
        ro, co = int(torch.randint(0, n - r, (1,), generator=g)), int(torch.randint(0, n - c, (1,), generator=g))
        win = torch.empty(r * c, dtype=torch.float32, device=cuda)
        rb.fill_dense(lay, D, r, c, ro, co, win, rb.RNGState(11))
        W = win.view(r, c) if lay == rb.Layout.RowMajor else win.view(c, r).t()
        assert torch.equal(W, F[ro:ro + r, co:co + c]), (r, c, ro, co)
    chk = full.view(torch.int32).view(-1, 1 << 16).sum(dim=1, dtype=torch.int64).sum()
    mean, var = float(full[: 1 << 28].double().mean()), float(full[: 1 << 28].double().var())
    assert abs(mean) < 1e-3 and abs(var - 1.0) < 1e-3
    full.zero_()
    rb.fill_dense(D, full, rb.RNGState(11))
    assert int(full.view(torch.int32).view(-1, 1 << 16).sum(dim=1, dtype=torch.int64).sum()) == int(chk)


def test_c4_sketch_sparse_full_size(rb, cuda):
    """C4: sketch_sparse, Gaussian d = 512, CSR 10^6 x 10^6 with 100 nonzeros per row (left_spmm at full size): columns of
    the result recomputed from the filled operator and the nonzeros of that column of A."""
    need_gb(cuda, 40)
    n, d, per_row = 1000000, 512, 100
    nnz = n * per_row
    cols = torch.empty(nnz, dtype=torch.int64, device=cuda)
    rb.repeated_fisher_yates(rb.RNGState(7), per_row, n, n, cols)
    cols = cols.view(n, per_row).sort(dim=1).values.reshape(-1).contiguous()
    vals = torch.empty(nnz, dtype=torch.float64, device=cuda)
    rb.fill_dense(rb.DenseDist(nnz, 1), vals, rb.RNGState(8))
    rowptr = torch.arange(0, nnz + 1, per_row, dtype=torch.int64, device=cuda)
    A = rb.CSRMatrix(n, n, vals, rowptr, cols)
    S = rb.DenseSkOp(rb.DenseDist(d, n), 3, dtype=torch.float64)
    B = torch.full((d * n,), float("nan"), dtype=torch.float64, device=cuda)
    L, O = rb.Layout.ColMajor, rb.Op.NoTrans
    rb.sketch_sparse(L, O, O, d, n, n, 1.0, S, 0, 0, A, 0, 0, 0.0, B, d)
    Bm = B.view(n, d)                                                      # Bm[j] = column j of B
    Sbuf = torch.empty(d * n, dtype=torch.float64, device=cuda)
    rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(d, n), d, n, 0, 0, Sbuf, rb.RNGState(3))
    Sm = Sbuf.view(n, d)                                                   # Sm[l] = column l of S
    rows_of = torch.arange(n, device=cuda).repeat_interleave(per_row)
    g = torch.Generator(device="cpu").manual_seed(9)
    for j in torch.randint(0, n, (12,), generator=g).tolist():
        sel = (cols == j).nonzero().squeeze(1)
        ref = (vals[sel].unsqueeze(1) * Sm[rows_of[sel]]).sum(dim=0)
        mag = (vals[sel].abs().unsqueeze(1) * Sm[rows_of[sel]].abs()).sum(dim=0) + 1e-300
        # the operator slab is held in float on the device (exact: its entries are floats) and accumulated in double
        assert float(((Bm[j] - ref).abs() / mag).max()) <= max(int(sel.numel()), 1) * 2.0 ** -52, j
    assert bool(torch.isfinite(B).all())
    B2 = torch.empty_like(B)
    rb.sketch_sparse(L, O, O, d, n, n, 1.0, S, 0, 0, A, 0, 0, 0.0, B2, d)
    assert torch.equal(B, B2)


def test_c5_column_slab_full_depth(rb, cuda):
    """C5: double, d = 8192, A 2^23 x 4096: the 512-column slab one GPU of the 8-GPU column-sharded configuration owns, at
    the full contraction length 2^23. Rows of B against S[i, :] A in FP64 (S row = a 1 x 2^23 window of the fill)."""
    need_gb(cuda, 60)
    d, m, n = 8192, 1 << 23, 512
    A = torch.empty(m * n, dtype=torch.float64, device=cuda)
    rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(m, 4096), m, n, 0, 3 * n, A, rb.RNGState(99))
    S = rb.DenseSkOp(rb.DenseDist(d, m), 5, dtype=torch.float64)
    B = torch.full((d * n,), float("nan"), dtype=torch.float64, device=cuda)
    L, O = rb.Layout.ColMajor, rb.Op.NoTrans
    rb.sketch_general(L, O, O, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d)
    Am = A.view(n, m)                                                      # Am[j] = column j of A
    Bm = B.view(n, d)
    g = torch.Generator(device="cpu").manual_seed(2)
    srow = torch.empty(m, dtype=torch.float64, device=cuda)
    for i in torch.randint(0, d, (6,), generator=g).tolist():
        rb.fill_dense(rb.Layout.ColMajor, rb.DenseDist(d, m), 1, m, i, 0, srow, rb.RNGState(5))
        ref = Am @ srow
        mag = Am.abs() @ srow.abs()
        assert float(((Bm[:, i] - ref).abs() / mag).max()) <= 32 * 2.0 ** -52, i
    assert bool(torch.isfinite(B).all())
