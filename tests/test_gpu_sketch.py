"""GPU parity tests for the sketching kernels (K2 dense, K4 sparse operator, K5 sparse data) through the host mirror
of the RandBLAS API. Tolerance = the reference's own componentwise bound (test/test_matmul_cores/linop_common.hh:257-263):
    |B_hat - B| <= |alpha| * K * 2*eps * (|S||A|) + |beta| * eps * |B0|
with eps the unit roundoff of T. Cases mirror test_lskge3.cc / test_rskge3.cc / test_lskges.cc / test_rskges.cc /
test_spmm_*.cc / test_sketch_vector.cc / test_sketch_symmetric.cc."""
import numpy as np
import pytest
import torch

import oracle_lib as ol

pytestmark = pytest.mark.gpu
TDT = {"f32": torch.float32, "f64": torch.float64}
EPS = {"f32": 2.0 ** -23, "f64": 2.0 ** -52}


def mat(layout, rows, cols, ld, data):
    """(rows x cols) numpy view of a flat buffer in `layout` with leading dimension ld"""
    if layout == "C":
        return np.lib.stride_tricks.as_strided(data, (rows, cols), (data.itemsize, ld * data.itemsize))
    return np.lib.stride_tricks.as_strided(data, (rows, cols), (ld * data.itemsize, data.itemsize))


def span(layout, rows, cols, ld):
    return ((cols - 1) * ld + rows) if layout == "C" else ((rows - 1) * ld + cols)


def dense_S(dt, D_rows, D_cols, fam, ma, ctr, key, rows, cols, ro, co):
    rc, out, _ = ol.fill_dense(dt, "R", D_rows, D_cols, fam, ma, rows, cols, ro, co, ctr, key)
    assert rc == 0
    return out.reshape(rows, cols).astype(np.float64)


def check_bound(dt, got, B0, Sop, Aop, alpha, beta, side, what):
    """got, B0: (rowsB x colsB) float64 arrays; Sop/Aop: explicit op(submat(S)), op(A) as float64"""
    P = Sop @ Aop if side == "L" else Aop @ Sop
    exact = alpha * P + (beta * B0 if beta != 0 else 0)
    K = Sop.shape[1] if side == "L" else Sop.shape[0]
    absP = np.abs(Sop) @ np.abs(Aop) if side == "L" else np.abs(Aop) @ np.abs(Sop)
    bound = abs(alpha) * K * 2 * EPS[dt] * absP + abs(beta) * EPS[dt] * np.abs(B0) + 1e-300
    ratio = float(np.max(np.abs(got - exact) / bound))
    assert ratio <= 1.0, f"{what}: error/bound = {ratio}"
    return ratio


def run_dense_case(rb, cuda, dt, side, layout, opS, opA, d, n, m, alpha, beta, D_extra=(0, 0), off=(0, 0), fam="G", ma="L",
                   lda_pad=0, ldb_pad=0, key=(3, 0), prefilled=False, host=False):
    """host=True: A and B stay in host memory (the rbx_host_* entry points stage them)"""
    rng = np.random.default_rng(d * 131 + n * 17 + m)
    npdt = np.float32 if dt == "f32" else np.float64
    if side == "L":
        rowsS, colsS = (d, m) if opS == "N" else (m, d)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB = d, n
        dims = (d, n, m)
    else:
        # B (m x d) = op(A) (m x n) op(S) (n x d)
        rowsS, colsS = (n, d) if opS == "N" else (d, n)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB = m, d
        dims = (m, d, n)
    Dr, Dc = rowsS + off[0] + D_extra[0], colsS + off[1] + D_extra[1]
    lda = (rowsA if layout == "C" else colsA) + lda_pad
    ldb = (rowsB if layout == "C" else colsB) + ldb_pad
    A = rng.standard_normal(span(layout, rowsA, colsA, lda)).astype(npdt)
    B0 = rng.standard_normal(span(layout, rowsB, colsB, ldb)).astype(npdt)
    S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), rb.RNGState(list(key)), dtype=TDT[dt])
    if prefilled:
        rb.fill_dense(S)
    where = "cpu" if host else cuda
    dA, dB = torch.from_numpy(A).to(where), torch.from_numpy(B0.copy()).to(where)
    L, O = rb.Layout(layout), rb.Op
    if side == "L":
        rb.sketch_general(L, O(opS), O(opA), *dims, alpha, S, off[0], off[1], dA, lda, beta, dB, ldb)
    else:
        rb.sketch_general(L, O(opA), O(opS), *dims, alpha, dA, lda, S, off[0], off[1], beta, dB, ldb)
    got_flat = dB.cpu().numpy()
    Sx = dense_S(dt, Dr, Dc, fam, ma, [0, 0, 0, 0], list(key), rowsS, colsS, off[0], off[1])
    if fam == "U" and dt == "f32":
        pass   # oracle already scales in float
    Sop = Sx if opS == "N" else Sx.T
    Aop = mat(layout, rowsA, colsA, lda, A).astype(np.float64)
    Aop = Aop if opA == "N" else Aop.T
    got = mat(layout, rowsB, colsB, ldb, got_flat).astype(np.float64)
    B0m = mat(layout, rowsB, colsB, ldb, B0).astype(np.float64)
    ratio = check_bound(dt, got, B0m, Sop, Aop, alpha, beta, side, (dt, side, layout, opS, opA, dims))
    # padding of B (ld > rows/cols) must be untouched
    if ldb_pad:
        mask = np.ones(len(B0), bool)
        idx = mat(layout, rowsB, colsB, ldb, np.arange(len(B0)))
        mask[idx.ravel()] = False
        assert np.array_equal(got_flat[mask], B0[mask])
    return ratio


@pytest.mark.parametrize("dt", ["f32", "f64"])
@pytest.mark.parametrize("layout", ["C", "R"])
@pytest.mark.parametrize("side", ["L", "R"])
def test_sketch_general_dense_small(rb, cuda, dt, layout, side):
    """test_lskge3.cc / test_rskge3.cc: sketch, lift, transposed S, submatrix of S (offsets (3,1) in 8x12), submatrix of A"""
    for opS in "NT":
        for opA in "NT":
            run_dense_case(rb, cuda, dt, side, layout, opS, opA, 7, 5, 19, 1.0, 0.0)
            run_dense_case(rb, cuda, dt, side, layout, opS, opA, 19, 4, 7, 1.0, 0.0)              # lift
    run_dense_case(rb, cuda, dt, side, layout, "N", "N", 5, 6, 11, 1.0, 0.0, D_extra=(0, 0), off=(3, 1))
    run_dense_case(rb, cuda, dt, side, layout, "T", "N", 5, 6, 11, 1.0, 0.0, D_extra=(2, 3), off=(3, 1))
    run_dense_case(rb, cuda, dt, side, layout, "N", "N", 7, 5, 19, 1.0, 0.0, lda_pad=3, ldb_pad=2)  # submatrix of A via lda
    run_dense_case(rb, cuda, dt, side, layout, "N", "T", 7, 5, 19, 5.5, -1.0, lda_pad=1)            # alpha / beta
    run_dense_case(rb, cuda, dt, side, layout, "N", "N", 7, 5, 19, 0.0, 2.0)
    for fam in "GU":
        for ma in "LS":
            run_dense_case(rb, cuda, dt, side, layout, "N", "N", 9, 3, 33, 1.0, 0.0, fam=fam, ma=ma, off=(1, 2))
            run_dense_case(rb, cuda, dt, side, layout, "T", "N", 9, 3, 33, 1.0, 0.0, fam=fam, ma=ma, off=(2, 1))
    run_dense_case(rb, cuda, dt, side, layout, "N", "N", 7, 5, 19, 1.0, 0.5, prefilled=True, off=(1, 1), D_extra=(1, 1))
    run_dense_case(rb, cuda, dt, side, layout, "T", "N", 7, 5, 19, 1.0, 0.5, prefilled=True)


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sketch_general_dense_medium_ragged(rb, cuda, dt):
    """sizes that are not multiples of any tile, both sides, unaligned operator offsets"""
    for (side, layout, opS, opA, d, n, m, off) in [("L", "C", "N", "N", 130, 70, 517, (0, 0)), ("L", "R", "N", "N", 65, 129, 300, (5, 3)),
                                                  ("L", "C", "T", "T", 100, 33, 257, (2, 7)), ("R", "C", "N", "N", 90, 64, 400, (1, 1)),
                                                  ("R", "R", "T", "N", 64, 128, 256, (0, 0)), ("L", "C", "N", "N", 1, 1, 1000, (0, 1))]:
        run_dense_case(rb, cuda, dt, side, layout, opS, opA, d, n, m, 1.0, 0.0, off=off, D_extra=(3, 5))


def test_sketch_general_dense_tensor_core_path(rb, cuda, monkeypatch):
    """float problems inside the tcgen05 kernel's envelope (rbx_sketch_dense_tc.cu): ragged M / N / K, operator offsets
    (aligned and not), alpha / beta, both families and major axes, the right-sketch form, a pre-filled operator. The launch
    count proves the path taken: fused kernel + split-K reduce = 2 launches (uniform family); k-panel form = panels x (panel
    fill + kernel + reduce) (Gaussian family, or an operator orientation the fused generator does not cover); the SIMT
    kernel would be 1."""
    cases = [("L", "C", "N", "N", 256, 300, 4100, 1.0, 0.0, (0, 0), "G", "L", 2),           # two column tiles: fused
             ("L", "C", "N", "N", 200, 600, 4096, 1.5, -0.5, (5, 8), "G", "L", 3),          # three: k-panel form
             ("L", "C", "N", "N", 200, 260, 4100, 1.5, -0.5, (5, 7), "G", "L", 3 * 4 + 2),  # unaligned offset; 4 panels of 1024 + one of 4 (no split-K)
             ("L", "C", "N", "N", 128, 512, 2052, 1.0, 0.0, (0, 4), "U", "L", 2),
             ("L", "C", "N", "N", 128, 512, 2052, 1.0, 0.5, (2, 1), "U", "S", 3),           # parent rows along i: panel form
             ("R", "R", "N", "N", 256, 300, 4100, 1.0, 2.0, (0, 0), "G", "L", 2),
             ("L", "R", "N", "T", 256, 256, 4096, 1.0, 0.0, (0, 0), "U", "L", 2)]
    for (side, layout, opS, opA, d, n, m, alpha, beta, off, fam, ma, want) in cases:
        if want > 3:
            monkeypatch.setenv("RBX_TC_PANEL_K", "1024")
        else:
            monkeypatch.delenv("RBX_TC_PANEL_K", raising=False)
        l0 = rb.lib.rbx_launch_count()
        ro, co = off
        if side == "R":
            # B (n x d) = A (n x m) * S (m x d): operator offsets are (rows, cols) of the tall S; its parent is S^T
            ratio = run_dense_case(rb, cuda, "f32", side, layout, opS, opA, d, m, n, alpha, beta, off=(co, ro), D_extra=(8, 3), fam=fam, ma=ma)
        else:
            ratio = run_dense_case(rb, cuda, "f32", side, layout, opS, opA, d, n, m, alpha, beta, off=(ro, co), D_extra=(3, 8), fam=fam, ma=ma)
        launches = rb.lib.rbx_launch_count() - l0
        assert launches == want, f"expected {want} launches on the tensor-core path, saw {launches} for {(side, layout, d, n, m, fam, ma)}"
        assert ratio < 0.05      # 3xTF32 with FP32 accumulation sits far inside the FP32 GEMM bound
    monkeypatch.delenv("RBX_TC_PANEL_K", raising=False)
    # both forms can be forced (RBX_TC_PANEL) and agree
    for force, want in (("0", 2), ("1", 3)):
        monkeypatch.setenv("RBX_TC_PANEL", force)
        l0 = rb.lib.rbx_launch_count()
        assert run_dense_case(rb, cuda, "f32", "L", "C", "N", "N", 256, 600, 4100, 1.0, 0.0, D_extra=(3, 8), fam="G") < 0.05
        assert rb.lib.rbx_launch_count() - l0 == want
    monkeypatch.delenv("RBX_TC_PANEL", raising=False)
    # DenseSkOp::buff given (k contiguous): fill_dense(S) + kernel + reduce
    l0 = rb.lib.rbx_launch_count()
    assert run_dense_case(rb, cuda, "f32", "L", "C", "N", "N", 192, 256, 4096, 1.0, 0.5, prefilled=True) < 0.05
    assert rb.lib.rbx_launch_count() - l0 == 3
    # data whose unit stride is NOT along the contraction axis (RowMajor A in a left sketch, ColMajor A in a right sketch, a
    # transposed ColMajor A): copied chunk by chunk into a k-contiguous workspace, then the same kernels. One chunk: copy +
    # kernel + reduce = 3 (uniform, fused) / copy + panel fill + kernel + reduce = 4 (Gaussian, three column tiles); with
    # RBX_TC_COPY_MB = 2 the 4100 x 300 data goes in 3 chunks of 1376 rows
    for (side, layout, opS, opA, d, n, m, alpha, beta, off, fam, ma, want, copy_mb) in [
            ("L", "R", "N", "N", 256, 300, 4100, 1.5, -0.5, (5, 8), "U", "L", 3, None),
            ("L", "R", "N", "N", 200, 600, 4096, 1.0, 0.0, (0, 0), "G", "L", 4, None),
            ("R", "C", "N", "N", 256, 300, 4100, 1.0, 2.0, (0, 0), "U", "L", 3, None),
            ("L", "C", "N", "T", 128, 512, 2052, 1.0, 0.5, (2, 4), "U", "L", 3, None),
            ("L", "R", "N", "N", 256, 300, 4100, 1.5, -0.5, (5, 8), "U", "L", 3 * 3, "2"),
            ("L", "C", "N", "N", 256, 300, 4100, 1.0, 0.0, (0, 0), "U", "L", 3, None)]:      # k-contiguous but lda = 4103 is not a multiple of 4: repacked
        if copy_mb:
            monkeypatch.setenv("RBX_TC_COPY_MB", copy_mb)
        l0 = rb.lib.rbx_launch_count()
        ro, co = off
        if side == "R":
            ratio = run_dense_case(rb, cuda, "f32", side, layout, opS, opA, d, m, n, alpha, beta, off=(co, ro), D_extra=(8, 3), fam=fam, ma=ma)
        else:
            ratio = run_dense_case(rb, cuda, "f32", side, layout, opS, opA, d, n, m, alpha, beta, off=(ro, co), D_extra=(3, 8), fam=fam, ma=ma, lda_pad=3)
        launches = rb.lib.rbx_launch_count() - l0
        monkeypatch.delenv("RBX_TC_COPY_MB", raising=False)
        assert launches == want, f"expected {want} launches (k-major copy route), saw {launches} for {(side, layout, opA, d, n, m, fam)}"
        assert ratio < 0.05


def test_sketch_general_dense_copy_route_by_column_blocks(rb, cuda, monkeypatch):
    """row-contiguous data with many columns: the k-major copy goes by column blocks that hold the whole contraction axis, each
    block of C written once. RBX_TC_COPY_MB = 4 makes the blocks 2048 columns wide at m = 512: 5000 columns = two full blocks on
    the tcgen05 kernel and a last one of 904 columns that falls below the work threshold (SIMT kernel on the copied block)."""
    monkeypatch.setenv("RBX_TC_COPY_MB", "4")
    for (side, layout, opA, fam, alpha, beta, ldb_pad) in [("L", "R", "N", "U", 1.5, -0.5, 0), ("L", "R", "N", "G", 1.0, 0.0, 4), ("L", "C", "T", "U", 1.0, 2.0, 0),
                                                          ("R", "C", "N", "G", 1.0, 0.5, 0)]:
        l0 = rb.lib.rbx_launch_count()
        if side == "L":
            ratio = run_dense_case(rb, cuda, "f32", side, layout, "N", opA, 128, 5000, 512, alpha, beta, off=(2, 4), D_extra=(1, 8), fam=fam, ldb_pad=ldb_pad)
        else:
            ratio = run_dense_case(rb, cuda, "f32", side, layout, "N", opA, 128, 512, 5000, alpha, beta, off=(4, 2), D_extra=(8, 1), fam=fam)
        assert ratio < 0.05
        assert rb.lib.rbx_launch_count() - l0 >= 7, (side, layout, fam, rb.lib.rbx_launch_count() - l0)
    monkeypatch.delenv("RBX_TC_COPY_MB")


def test_sketch_general_dense_dmma_path(rb, cuda, monkeypatch):
    """double problems inside the DMMA kernel's envelope (rbx_sketch_dense_dmma.cu): ragged M / N / K (odd K included),
    operator offsets, alpha / beta, both families and major axes, both data layouts, the right-sketch form, a pre-filled
    operator, and several k-panels (RBX_DMMA_PANEL shrinks the panel). Launches = panels x (operator panel fill + DMMA
    GEMM); the SIMT kernel would be 1."""
    cases = [("L", "C", "N", "N", 200, 260, 4100, 1.0, 0.0, (0, 0), "G", "L", 0, 1),
             ("L", "C", "N", "N", 200, 260, 4101, 1.5, -0.5, (5, 8), "G", "L", 1, 5),
             ("L", "R", "N", "N", 130, 260, 4100, 1.0, 0.0, (0, 3), "U", "L", 0, 5),
             ("L", "C", "N", "N", 128, 512, 2052, 1.0, 0.5, (2, 1), "G", "S", 0, 3),
             ("R", "R", "N", "N", 256, 300, 4100, 1.0, 2.0, (0, 0), "G", "L", 0, 1),
             ("L", "R", "N", "T", 256, 256, 4096, 1.0, 0.0, (0, 0), "U", "L", 0, 4)]
    for (side, layout, opS, opA, d, n, m, alpha, beta, off, fam, ma, lda_pad, panels) in cases:
        if panels > 1:
            monkeypatch.setenv("RBX_DMMA_PANEL", "1024")
        else:
            monkeypatch.delenv("RBX_DMMA_PANEL", raising=False)
        l0 = rb.lib.rbx_launch_count()
        ro, co = off
        if side == "R":
            ratio = run_dense_case(rb, cuda, "f64", side, layout, opS, opA, d, m, n, alpha, beta, off=(co, ro), D_extra=(8, 3), fam=fam, ma=ma)
        else:
            ratio = run_dense_case(rb, cuda, "f64", side, layout, opS, opA, d, n, m, alpha, beta, off=(ro, co), D_extra=(3, 8), fam=fam, ma=ma,
                                   lda_pad=lda_pad)
        launches = rb.lib.rbx_launch_count() - l0
        # per panel: operator fill + DMMA GEMM (+ the split-K reduce: these shapes have few output tiles)
        assert 2 * panels <= launches <= 3 * panels, f"expected {panels} panel(s) of fill + DMMA GEMM (+ reduce), saw {launches} launch(es) for {(side, layout, d, n, m)}"
        assert ratio < 0.05
    monkeypatch.delenv("RBX_DMMA_PANEL", raising=False)
    # DenseSkOp::buff given (explicit operator, k contiguous): one DMMA launch, no fill
    l0 = rb.lib.rbx_launch_count()
    run_dense_case(rb, cuda, "f64", "L", "C", "N", "N", 192, 256, 4096, 1.0, 0.5, prefilled=True)
    assert rb.lib.rbx_launch_count() - l0 in (2, 3)      # fill_dense(S) + DMMA GEMM (+ split-K reduce)
    # tall-skinny data (few columns): tensor-core kernels with split-K in both precisions instead of the SIMT kernel
    for dt, want in (("f32", (2, 3)), ("f64", (3,))):
        for (layout, d, n, m) in (("C", 256, 5, 120000), ("R", 192, 17, 45000), ("C", 128, 2, 600000)):
            l0 = rb.lib.rbx_launch_count()
            assert run_dense_case(rb, cuda, dt, "L", layout, "N", "N", d, n, m, 1.5, -0.5, fam="G" if n != 17 else "U", lda_pad=0 if layout == "C" else 3) < 0.05
            got = rb.lib.rbx_launch_count() - l0
            assert got >= 2 and got != 1, (dt, layout, d, n, m, got)
        # a small sketch dimension (d = 1 .. 48): the operator tile is mostly padding, the data still streams through the tensor-core kernels
        for (layout, d, n, m) in (("C", 5, 300, 100000), ("R", 33, 200, 30000), ("C", 1, 1000, 150000)):
            l0 = rb.lib.rbx_launch_count()
            assert run_dense_case(rb, cuda, dt, "L", layout, "N", "N", d, n, m, 1.0, 0.0, fam="G", off=(0, 4), D_extra=(2, 8)) < 0.05
            assert rb.lib.rbx_launch_count() - l0 >= 2, (dt, layout, d, n, m)


def test_dense_tensor_core_paths_take_misaligned_operands(rb, cuda):
    """operands the 16-byte copies of the tensor-core kernels cannot read as they are still run on those kernels, not on the
    SIMT fallback (1 launch): double data with an odd leading dimension in either layout (8-byte staging copies in the DMMA
    kernel); a pre-filled operator with an odd row length or a window offset that breaks the alignment (DMMA: 8-byte copies;
    tcgen05: repacked panel); a pre-filled short-axis-major operator, whose unit stride runs along the sketch dimension
    (transposed panel by panel into a k-contiguous workspace in both precisions)."""
    def launches(**kw):
        l0 = rb.lib.rbx_launch_count()
        assert run_dense_case(rb, cuda, **kw) < 0.05
        return rb.lib.rbx_launch_count() - l0
    base = dict(side="L", opS="N", opA="N", alpha=1.5, beta=-0.5)
    # implicit operator, double, odd lda: operator panel fill + DMMA GEMM (+ split-K reduce)
    assert launches(dt="f64", layout="C", d=200, n=260, m=4101, **base) >= 2
    assert launches(dt="f64", layout="R", d=130, n=261, m=4100, off=(3, 1), D_extra=(2, 2), **base) >= 2
    assert launches(dt="f64", layout="C", d=200, n=260, m=4100, lda_pad=3, **base) >= 2
    # pre-filled operators: fill_dense(S) is one launch and the SIMT kernel one more
    for dt in ("f32", "f64"):
        assert launches(dt=dt, layout="C", d=192, n=256, m=4099, prefilled=True, **base) >= 3                       # odd row length
        assert launches(dt=dt, layout="C", d=192, n=256, m=4096, prefilled=True, off=(1, 3), D_extra=(2, 2), **base) >= 3  # window offset
        assert launches(dt=dt, layout="C", d=192, n=256, m=4096, prefilled=True, ma="S", **base) >= 3               # unit stride along i
        assert launches(dt=dt, layout="R", d=190, n=300, m=4101, prefilled=True, ma="S", off=(2, 5), D_extra=(1, 1), **base) >= 3
        assert launches(dt=dt, side="R", layout="R", opS="N", opA="N", alpha=1.0, beta=2.0, d=256, n=4100, m=300, prefilled=True, ma="S") >= 3


def test_nan_in_B_is_overwritten_when_beta_is_zero(rb, cuda):
    """util::safe_scal (util.hh:51-59) / BLAS convention: beta == 0 never reads B"""
    S = rb.DenseSkOp(rb.DenseDist(4, 9), 1)
    A = torch.ones(9 * 3, device=cuda)
    B = torch.full((12,), float("nan"), device=cuda)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 4, 3, 9, 1.0, S, 0, 0, A, 9, 0.0, B, 4)
    assert bool(torch.isfinite(B).all())
    Ss = rb.SparseSkOp(rb.SparseDist(4, 9, 2), 1, dtype=torch.float32)
    B = torch.full((12,), float("nan"), device=cuda)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 4, 3, 9, 1.0, Ss, 0, 0, A, 9, 0.0, B, 4)
    assert bool(torch.isfinite(B).all())


# ---------------------------------------------------------------------------------------------- sparse operators
def sparse_S_dense(dt, r, c, k, ma, ctr, key):
    rc, rows, cols, vals, _ = ol.fill_sparse("f64", "i64", r, c, k, ma, ctr, key)
    assert rc == 0
    S = np.zeros((r, c))
    np.add.at(S, (rows, cols), vals)
    return S


def run_sparseop_case(rb, cuda, dt, side, layout, opS, opA, d, n, m, alpha, beta, k, ma, D_extra=(0, 0), off=(0, 0), key=42,
                      lda_pad=0, filled=False, host=False):
    rng = np.random.default_rng(d * 7 + n * 3 + m + k)
    npdt = np.float32 if dt == "f32" else np.float64
    if side == "L":
        rowsS, colsS = (d, m) if opS == "N" else (m, d)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB, dims = d, n, (d, n, m)
    else:
        rowsS, colsS = (n, d) if opS == "N" else (d, n)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB, dims = m, d, (m, d, n)
    Sr, Sc = rowsS + off[0] + D_extra[0], colsS + off[1] + D_extra[1]
    lda = (rowsA if layout == "C" else colsA) + lda_pad
    ldb = rowsB if layout == "C" else colsB
    A = rng.standard_normal(span(layout, rowsA, colsA, lda)).astype(npdt)
    B0 = rng.standard_normal(rowsB * colsB).astype(npdt)
    S = rb.SparseSkOp(rb.SparseDist(Sr, Sc, k, ma), key, dtype=TDT[dt])
    if filled:
        rb.fill_sparse(S)
    where = "cpu" if host else cuda
    dA, dB = torch.from_numpy(A).to(where), torch.from_numpy(B0.copy()).to(where)
    L, O = rb.Layout(layout), rb.Op
    if side == "L":
        rb.sketch_general(L, O(opS), O(opA), *dims, alpha, S, off[0], off[1], dA, lda, beta, dB, ldb)
    else:
        rb.sketch_general(L, O(opA), O(opS), *dims, alpha, dA, lda, S, off[0], off[1], beta, dB, ldb)
    Sfull = sparse_S_dense(dt, Sr, Sc, k, ma, [0, 0, 0, 0], [key, 0])
    Sx = Sfull[off[0]:off[0] + rowsS, off[1]:off[1] + colsS]
    Sop = Sx if opS == "N" else Sx.T
    Aop = mat(layout, rowsA, colsA, lda, A).astype(np.float64)
    Aop = Aop if opA == "N" else Aop.T
    got = mat(layout, rowsB, colsB, ldb, dB.cpu().numpy()).astype(np.float64)
    B0m = mat(layout, rowsB, colsB, ldb, B0).astype(np.float64)
    check_bound(dt, got, B0m, Sop, Aop, alpha, beta, side, ("sparse", dt, side, layout, opS, opA, dims, k, ma))


@pytest.mark.parametrize("dt", ["f32", "f64"])
@pytest.mark.parametrize("layout", ["C", "R"])
def test_sketch_general_sparse_operator(rb, cuda, dt, layout):
    """test_lskges.cc:174-589 / test_rskges.cc: SASO & LASO, keys {42,0,1}, vec_nnz {1,2,3,7,19}, 19x201 * 201x12 and lifts,
    row/col subsets of S, transposes, submatrix of A, alpha/beta."""
    for ma in "SL":
        for key in (42, 0, 1):
            for k in (1, 2, 3, 7, 19):
                run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 19, 12, 201, 1.0, 0.0, k, ma, key=key)
        for k in (1, 3, 7):
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 201, 12, 19, 1.0, 0.0, k, ma)          # lift (tall S)
            run_sparseop_case(rb, cuda, dt, "R", layout, "N", "N", 19, 12, 201, 1.0, 0.0, k, ma)          # A (12x201) S (201x19)
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 12, 9, 150, 1.0, 0.0, k, ma, D_extra=(7, 51), off=(0, 0))
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 12, 9, 150, 1.0, 0.0, k, ma, D_extra=(3, 11), off=(4, 40))
            run_sparseop_case(rb, cuda, dt, "L", layout, "T", "N", 19, 12, 201, 1.0, 0.0, k, ma)
            run_sparseop_case(rb, cuda, dt, "R", layout, "T", "T", 19, 12, 201, 1.0, 0.0, k, ma, off=(1, 2), D_extra=(1, 1))
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "T", 19, 12, 201, 1.0, 0.0, k, ma, lda_pad=5)
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 19, 12, 201, 5.5, -1.0, k, ma)
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 19, 12, 201, 5.5, 0.0, k, ma, filled=True)
            run_sparseop_case(rb, cuda, dt, "R", layout, "N", "N", 19, 12, 201, 1.0, 0.5, k, ma, filled=True, off=(2, 1), D_extra=(0, 3))


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_saso_medium(rb, cuda, dt):
    """a config-2-shaped case at 1/64 scale in each dimension class: d=256, m=16384, n=96, vec_nnz=8, both layouts, offsets"""
    for layout in "CR":
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 256, 96, 16384, 1.0, 0.0, 8, "S", key=1997)
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 200, 100, 10001, 2.0, 1.0, 8, "S", key=1997, off=(5, 1234), D_extra=(51, 77))
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 2048, 64, 4096, 1.0, 0.0, 8, "S", key=7)
        # the usual right sketch B = A * S with a tall SASO S (and its left form with opS = Trans): same streaming kernel,
        # the operator being the wide one of the transposed shape with the same state
        run_sparseop_case(rb, cuda, dt, "R", layout, "N", "N", 256, 96, 16384, 1.0, 0.0, 8, "S", key=1997)
        run_sparseop_case(rb, cuda, dt, "R", layout, "N", "T", 200, 100, 10001, 2.0, 1.0, 8, "S", key=1997, off=(1234, 5), D_extra=(77, 51))
        run_sparseop_case(rb, cuda, dt, "L", layout, "T", "N", 256, 96, 16384, 1.5, 0.0, 8, "S", key=3)
    # data whose leading dimension is not a multiple of 16 bytes still takes the streaming kernel (element-wise cp.async needs
    # only the natural alignment): prep + tile + spill variant + reduce = 4 launches, not the general path's regroup
    l0 = rb.lib.rbx_launch_count()
    run_sparseop_case(rb, cuda, dt, "L", "C", "N", "N", 256, 96, 16384, 1.0, 0.5, 8, "S", key=11, lda_pad=1)
    assert rb.lib.rbx_launch_count() - l0 <= 4
    l0 = rb.lib.rbx_launch_count()
    run_sparseop_case(rb, cuda, dt, "L", "R", "N", "N", 256, 97, 16384, 1.0, 0.5, 8, "S", key=11, lda_pad=0)
    assert rb.lib.rbx_launch_count() - l0 <= 4
    l0 = rb.lib.rbx_launch_count()
    run_sparseop_case(rb, cuda, dt, "L", "R", "N", "N", 300, 40, 9000, 1.5, 0.0, 8, "S", key=12, lda_pad=3, off=(2, 77), D_extra=(5, 9))
    assert rb.lib.rbx_launch_count() - l0 <= 4


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_saso_small_windows_are_replicated(rb, cuda, dt, monkeypatch):
    """windows of at most 512 rows run the streaming kernel on R = 1024 // M copies of their rows (column cc of a chunk feeds copy
    cc mod R), summed by the reduce kernel: right against the oracle for M that do and do not divide 1024, with offsets and
    beta, equal to the unreplicated run to rounding, and bit-identical from run to run"""
    for layout in "CR":
        for (d, n, m, off, extra) in [(16, 50, 9000, (0, 0), (0, 0)), (100, 33, 12000, (3, 500), (7, 900)), (300, 64, 5000, (0, 0), (0, 0)), (512, 40, 4096, (0, 0), (0, 0))]:
            run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", d, n, m, 1.5, -0.5, 8, "S", key=21, off=off, D_extra=extra)
    npdt = np.float32 if dt == "f32" else np.float64
    rng = np.random.default_rng(3)
    d, n, m = 48, 96, 20000
    A = torch.from_numpy(rng.standard_normal(m * n).astype(npdt)).to(cuda)
    S = rb.SparseSkOp(rb.SparseDist(d, m, 8), 5, dtype=TDT[dt])
    outs = []
    for mode in ("repl", "repl", "norepl"):
        if mode == "norepl":
            monkeypatch.setenv("RBX_SASO_NO_REPL", "1")
        B = torch.zeros(d * n, dtype=TDT[dt], device=cuda)
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d)
        outs.append(B)
    monkeypatch.delenv("RBX_SASO_NO_REPL", raising=False)
    assert torch.equal(outs[0], outs[1])
    tol = 1e-4 if dt == "f32" else 1e-12
    assert torch.allclose(outs[0], outs[2], rtol=tol, atol=tol * float(outs[2].abs().max()))


def test_saso_ring_fills_shared_memory_to_the_byte(rb, cuda):
    """a draw of the randomized tests (block 186 of a 40x sweep): 15 x 554 window, vec_nnz = 24, 7 columns -- the ring stages came to
    exactly 227 KB of dynamic shared memory, which with the kernel's static mbarriers was one launch-time 'invalid argument'"""
    run_sparseop_case(rb, cuda, "f32", "R", "R", "T", "T", 15, 554, 7, -0.7, 1.0, 24, "S", D_extra=(7, 6), off=(2, 4), key=1523932692, lda_pad=1)


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_saso_tail_paths(rb, cuda, dt, monkeypatch):
    """The streaming SASO kernel's three homes of an entry: the slot record (first three per row and chunk), the staged
    tail region, and -- when a bucket's tails do not fit the region -- global memory. RBX_SASO_CAP shrinks the region so
    that the last path runs; vec_nnz = 32 on a short operator makes every row's tail long."""
    cases = [("L", "C", 256, 96, 16384, 8), ("L", "R", 200, 40, 5000, 8), ("L", "R", 64, 48, 2048, 32), ("L", "C", 2048, 32, 4096, 8),
             ("L", "R", 130, 17, 777, 19)]
    for cap in (None, "0", "32", "96"):
        if cap is None:
            monkeypatch.delenv("RBX_SASO_CAP", raising=False)
        else:
            monkeypatch.setenv("RBX_SASO_CAP", cap)
        for side, layout, d, n, m, k in cases:
            run_sparseop_case(rb, cuda, dt, side, layout, "N", "N", d, n, m, 1.5, 0.0, k, "S", key=3)
    monkeypatch.delenv("RBX_SASO_CAP", raising=False)


@pytest.mark.parametrize("fmt", ["O", "R", "C"])
@pytest.mark.parametrize("dc", ["8", "16", "32"])
def test_sketch_sparse_row_blocked_passes(rb, cuda, fmt, dc, monkeypatch):
    """the L2-resident schedule of sketch_sparse (spdata_gather_narrow_kernel: dc output rows per pass, operator slab of
    dc x K floats, slabs generated on the auxiliary stream while the previous pass gathers), forced at small sizes with
    RBX_SPD_DC: every format, both sides and layouts, both operator orientations and families, M not a multiple of dc,
    alpha / beta, and bitwise reproducibility. The full-size C4 test takes this schedule by default."""
    monkeypatch.setenv("RBX_SPD_DC", dc)
    run_sparsedata_case(rb, cuda, "f64", "L", "C", "N", "N", 601, 300, 700, 1.5, -0.5, fmt, 0.05, off_s=(1, 2))
    run_sparsedata_case(rb, cuda, "f32", "R", "R", "N", "N", 130, 400, 500, 1.0, 0.0, fmt, 0.1, fam="U")
    run_sparsedata_case(rb, cuda, "f64", "L", "R", "T", "T", 64, 300, 700, 1.0, 2.0, fmt, 0.05, ma="S")
    run_sparsedata_case(rb, cuda, "f64", "R", "C", "N", "N", 37, 90, 250, 2.0, 0.0, fmt, 0.3, fam="U", ma="S")
    rng = np.random.default_rng(5)
    A = to_fmt(rb, cuda, random_sparse(rng, 2000, 1500, 0.02, np.float64), fmt)
    S = rb.DenseSkOp(rb.DenseDist(96, 2000), 3, dtype=torch.float64)
    outs = []
    for _ in range(3):
        B = torch.zeros(96 * 1500, dtype=torch.float64, device=cuda)
        rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 96, 1500, 2000, 1.0, S, 0, 0, A, 0, 0, 0.0, B, 96)
        outs.append(B.cpu().numpy())
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    monkeypatch.delenv("RBX_SPD_DC", raising=False)


@pytest.mark.parametrize("fmt", ["O", "R", "C"])
def test_left_and_right_spmm_public_entry_points(rb, cuda, fmt):
    """sparse_data::left_spmm / right_spmm (spmm_dispatch.hh:48-200; test_spmm_{coo,csr,csc}.cc) on device containers: both
    layouts, transposes of either operand, alpha / beta, sub-matrix offsets (COO), against a dense product"""
    rng = np.random.default_rng(21)
    d, n, m = 23, 17, 140
    for layout in "CR":
        for opA in "NT":
            for opB in "NT":
                Am = random_sparse(rng, *((d, m) if opA == "N" else (m, d)), 0.15, np.float64)
                Bm = rng.standard_normal((m, n) if opB == "N" else (n, m))
                C0 = rng.standard_normal((d, n))
                A = to_fmt(rb, cuda, Am, fmt)
                ldb = Bm.shape[0] if layout == "C" else Bm.shape[1]
                ldc = d if layout == "C" else n
                Bf = torch.from_numpy(np.ascontiguousarray(Bm.T if layout == "C" else Bm).ravel()).to(cuda)
                Cf = torch.from_numpy(np.ascontiguousarray(C0.T if layout == "C" else C0).ravel()).to(cuda)
                rb.left_spmm(rb.Layout(layout), rb.Op(opA), rb.Op(opB), d, n, m, 1.5, A, 0, 0, Bf, ldb, -0.5, Cf, ldc)
                want = 1.5 * (Am if opA == "N" else Am.T) @ (Bm if opB == "N" else Bm.T) - 0.5 * C0
                got = Cf.cpu().numpy().reshape((n, d)).T if layout == "C" else Cf.cpu().numpy().reshape((d, n))
                assert np.allclose(got, want, rtol=1e-12, atol=1e-12), (fmt, layout, opA, opB)
    # right_spmm: C (m x d) = A (m x n dense) * B (n x d sparse)
    mm, dd, nn = 19, 11, 90
    Bs = random_sparse(rng, nn, dd, 0.2, np.float64)
    Ad = rng.standard_normal((mm, nn))
    C = torch.zeros(mm * dd, dtype=torch.float64, device=cuda)
    rb.right_spmm(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, mm, dd, nn, 1.0, torch.from_numpy(np.ascontiguousarray(Ad.T).ravel()).to(cuda), mm,
                  to_fmt(rb, cuda, Bs, fmt), 0, 0, 0.0, C, mm)
    assert np.allclose(C.cpu().numpy().reshape(dd, mm).T, Ad @ Bs, rtol=1e-12, atol=1e-12)
    if fmt == "O":        # sub-matrix of a COO operand
        Big = random_sparse(rng, 40, 160, 0.2, np.float64)
        A = to_fmt(rb, cuda, Big, "O")
        Bm = rng.standard_normal((m, n))
        Cf = torch.zeros(d * n, dtype=torch.float64, device=cuda)
        rb.left_spmm(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, A, 5, 9, torch.from_numpy(np.ascontiguousarray(Bm.T).ravel()).to(cuda), m, 0.0, Cf, d)
        assert np.allclose(Cf.cpu().numpy().reshape(n, d).T, Big[5:5 + d, 9:9 + m] @ Bm, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_general_sparse_operator_path_is_deterministic(rb, cuda, dt):
    """LASO, tall operators, explicit (filled) triplets and vec_nnz > 32 run the output-stationary gather of
    rbx_spmm_gather.cu (no atomics): correct against the oracle -- including rows long enough to be cut into sub-segments --
    and bit-identical from run to run, for both layouts of the data."""
    for layout in "CR":
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 24, 70, 20000, 1.5, -0.5, 6, "L", key=5)         # LASO, 5000 entries per row
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 64, 33, 6000, 1.0, 0.0, 40, "S", key=5)          # vec_nnz > 32
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 300, 140, 40, 1.0, 0.5, 3, "S", key=5)           # tall S (lift)
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 16, 130, 9000, 2.0, 0.0, 4, "S", key=5, filled=True)
    npdt = np.float32 if dt == "f32" else np.float64
    rng = np.random.default_rng(1)
    for (d, n, m, k, ma, layout) in [(24, 70, 20000, 6, "L", "C"), (24, 70, 20000, 6, "L", "R"), (64, 33, 6000, 40, "S", "R")]:
        A = torch.from_numpy(rng.standard_normal(m * n).astype(npdt)).to(cuda)
        S = rb.SparseSkOp(rb.SparseDist(d, m, k, ma), 11, dtype=TDT[dt])
        lda, ldb = (m, d) if layout == "C" else (n, n)
        outs = []
        for _ in range(3):
            B = torch.zeros(d * n, dtype=TDT[dt], device=cuda)
            rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, B, ldb)
            outs.append(B)
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2]), (d, n, m, k, ma, layout)


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_general_sparse_operator_path_relays_out_column_contiguous_data(rb, cuda, dt, monkeypatch):
    """column-contiguous data with >= 128 columns on the general path (LASO, vec_nnz > 32, explicit triplets): blocks of columns
    are re-laid with contiguous rows into a workspace and gathered with the row kernel (one extra launch per block, visible in
    the launch count), in one block and -- with RBX_CG_RELAYOUT_MB = 24 -- in several (identical bits), and the same values as
    without the re-layout to rounding; ragged column counts and a padded leading dimension included"""
    npdt = np.float32 if dt == "f32" else np.float64
    rng = np.random.default_rng(2)
    for (d, n, m, k, ma, lda_pad) in [(24, 300, 20000, 600, "L", 0), (64, 130, 6000, 40, "S", 3), (20, 257, 3000, 300, "L", 1)]:
        run_sparseop_case(rb, cuda, dt, "L", "C", "N", "N", d, n, m, 1.5, -0.5, k, ma, key=5, lda_pad=lda_pad)
        A = torch.from_numpy(rng.standard_normal((m + lda_pad) * n).astype(npdt)).to(cuda)
        S = rb.SparseSkOp(rb.SparseDist(d, m, k, ma), 11, dtype=TDT[dt])
        outs, launches = [], []
        for mode in ("default", "blocks", "off"):
            monkeypatch.delenv("RBX_CG_RELAYOUT_MB", raising=False); monkeypatch.delenv("RBX_CG_NO_RELAYOUT", raising=False)
            if mode == "blocks":
                monkeypatch.setenv("RBX_CG_RELAYOUT_MB", "24")
            if mode == "off":
                monkeypatch.setenv("RBX_CG_NO_RELAYOUT", "1")
            B = torch.full((d * n,), 0.25, dtype=TDT[dt], device=cuda)
            l0 = rb.lib.rbx_launch_count()
            rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 2.0, S, 0, 0, A, m + lda_pad, 0.5, B, d)
            launches.append(rb.lib.rbx_launch_count() - l0)
            outs.append(B)
        monkeypatch.delenv("RBX_CG_RELAYOUT_MB", raising=False); monkeypatch.delenv("RBX_CG_NO_RELAYOUT", raising=False)
        # one block or several: the same row kernel, the same order of additions -> identical bits; the 4-column kernel of the
        # route without re-layout folds its lanes' partial sums by a butterfly, i.e. another (equally fixed) order
        assert torch.equal(outs[0], outs[1]), (d, n, m, k, ma)
        tol = 1e-4 if dt == "f32" else 1e-12
        assert torch.allclose(outs[0], outs[2], rtol=tol, atol=tol * float(outs[2].abs().max())), (d, n, m, k, ma)
        assert launches[0] == launches[2] + 1 and launches[1] >= launches[0], (launches, d, n, m)


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_saso_windows_up_to_8192_rows_take_the_streaming_kernel(rb, cuda, dt):
    """operator windows of 4097..8192 rows (5 to 8 row blocks of the tile kernel): same result as the definition, and the launch
    count shows the streaming kernel ran (prep + two tile variants [+ reduce]), not fill_sparse + regroup + gather"""
    for layout in "CR":
        l0 = rb.lib.rbx_launch_count()
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 5000, 40, 12000, 1.0, 0.0, 8, "S", key=9)
        assert rb.lib.rbx_launch_count() - l0 <= 4
        run_sparseop_case(rb, cuda, dt, "L", layout, "N", "N", 8192, 24, 9000, 1.5, -1.0, 8, "S", key=9)


# ---------------------------------------------------------------------------------------------- sparse data
def random_sparse(rng, rows, cols, density, npdt):
    M = rng.standard_normal((rows, cols)) * (rng.random((rows, cols)) < density)
    return M.astype(npdt)


def to_fmt(rb, cuda, M, fmt):
    import scipy.sparse as sp
    t = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a if dt is None else a.astype(dt))).to(cuda)
    if fmt == "O":
        c = sp.coo_matrix(M)
        return rb.COOMatrix(M.shape[0], M.shape[1], t(c.data), t(c.row, np.int64), t(c.col, np.int64))
    if fmt == "R":
        c = sp.csr_matrix(M)
        return rb.CSRMatrix(M.shape[0], M.shape[1], t(c.data), t(c.indptr, np.int64), t(c.indices, np.int64))
    c = sp.csc_matrix(M)
    return rb.CSCMatrix(M.shape[0], M.shape[1], t(c.data), t(c.indices, np.int64), t(c.indptr, np.int64))


def run_sparsedata_case(rb, cuda, dt, side, layout, opS, opA, d, n, m, alpha, beta, fmt, density, off_s=(0, 0), off_a=(0, 0), fam="G", ma="L",
                        host=False):
    rng = np.random.default_rng(d + 3 * n + 7 * m + int(100 * density))
    npdt = np.float32 if dt == "f32" else np.float64
    if side == "L":
        rowsS, colsS = (d, m) if opS == "N" else (m, d)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB, dims = d, n, (d, n, m)
    else:
        rowsS, colsS = (n, d) if opS == "N" else (d, n)
        rowsA, colsA = (m, n) if opA == "N" else (n, m)
        rowsB, colsB, dims = m, d, (m, d, n)
    Dr, Dc = rowsS + off_s[0] + 2, colsS + off_s[1] + 1
    Ar, Ac = rowsA + off_a[0] + (3 if fmt == "O" and off_a != (0, 0) else 0), colsA + off_a[1]
    Afull = random_sparse(rng, Ar, Ac, density, npdt)
    where = "cpu" if host else cuda
    Asp = to_fmt(rb, where, Afull, fmt)
    ldb = rowsB if layout == "C" else colsB
    B0 = rng.standard_normal(rowsB * colsB).astype(npdt)
    dB = torch.from_numpy(B0.copy()).to(where)
    S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), 11, dtype=TDT[dt])
    L, O = rb.Layout(layout), rb.Op
    if side == "L":
        rb.sketch_sparse(L, O(opS), O(opA), *dims, alpha, S, off_s[0], off_s[1], Asp, off_a[0], off_a[1], beta, dB, ldb)
    else:
        rb.sketch_sparse(L, O(opA), O(opS), *dims, alpha, Asp, off_a[0], off_a[1], S, off_s[0], off_s[1], beta, dB, ldb)
    Sx = dense_S(dt, Dr, Dc, fam, ma, [0, 0, 0, 0], [11, 0], rowsS, colsS, off_s[0], off_s[1])
    Sop = Sx if opS == "N" else Sx.T
    Ax = Afull[off_a[0]:off_a[0] + rowsA, off_a[1]:off_a[1] + colsA].astype(np.float64)
    Aop = Ax if opA == "N" else Ax.T
    got = mat(layout, rowsB, colsB, ldb, dB.cpu().numpy()).astype(np.float64)
    B0m = mat(layout, rowsB, colsB, ldb, B0).astype(np.float64)
    check_bound(dt, got, B0m, Sop, Aop, alpha, beta, side, ("spdata", dt, side, layout, opS, opA, dims, fmt))


@pytest.mark.parametrize("dt", ["f32", "f64"])
@pytest.mark.parametrize("fmt", ["O", "R", "C"])
def test_sketch_sparse(rb, cuda, dt, fmt):
    """test_spmm_{coo,csr,csc}.cc through lsksp3/rsksp3: p in {0.01, 0.1, 0.8}, alpha/beta, transposes, submatrices
    (of the sparse matrix: COO only, spmm_dispatch.hh:99-104)"""
    for layout in "CR":
        for side in "LR":
            for density in (0.01, 0.1, 0.8):
                run_sparsedata_case(rb, cuda, dt, side, layout, "N", "N", 7, 15, 40, 1.0, 0.0, fmt, density)
            for opS in "NT":
                for opA in "NT":
                    run_sparsedata_case(rb, cuda, dt, side, layout, opS, opA, 6, 10, 25, 0.5, -1.5, fmt, 0.2, off_s=(1, 2))
            if fmt == "O":
                run_sparsedata_case(rb, cuda, dt, side, layout, "N", "N", 6, 10, 25, 1.0, 0.0, fmt, 0.3, off_a=(2, 3))
                run_sparsedata_case(rb, cuda, dt, side, layout, "N", "T", 6, 10, 25, 1.0, 0.0, fmt, 0.3, off_a=(1, 4))
            for fam in "GU":
                for ma in "LS":
                    run_sparsedata_case(rb, cuda, dt, side, layout, "N", "N", 9, 12, 31, 1.0, 0.0, fmt, 0.15, fam=fam, ma=ma, off_s=(2, 3))
    if fmt != "O":
        A = to_fmt(rb, cuda, random_sparse(np.random.default_rng(0), 20, 10, 0.3, np.float32 if dt == "f32" else np.float64), fmt)
        S = rb.DenseSkOp(rb.DenseDist(5, 20), 0, dtype=TDT[dt])
        B = torch.zeros(50, dtype=TDT[dt], device=cuda)
        with pytest.raises(rb.Error):        # CSR/CSC: whole matrix only
            rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 5, 9, 20, 1.0, S, 0, 0, A, 0, 1, 0.0, B, 5)


@pytest.mark.parametrize("fmt", ["O", "R", "C"])
def test_sketch_sparse_medium_slabs_and_determinism(rb, cuda, fmt, monkeypatch):
    """rbx_spdata.cu at sizes that use several 512-row chunks of the operator (M = 601: padded pitch), several warps per
    column pass and several contraction slabs (RBX_SPD_SLAB); the regrouped formats must give bitwise identical results
    from run to run (the column segments are rank-sorted after the atomic scatter)."""
    run_sparsedata_case(rb, cuda, "f64", "L", "C", "N", "N", 601, 300, 700, 1.5, -0.5, fmt, 0.05, off_s=(1, 2))
    run_sparsedata_case(rb, cuda, "f32", "R", "R", "N", "N", 130, 400, 500, 1.0, 0.0, fmt, 0.1, fam="U")
    monkeypatch.setenv("RBX_SPD_SLAB", "97")
    run_sparsedata_case(rb, cuda, "f64", "L", "C", "N", "N", 601, 300, 700, 1.5, -0.5, fmt, 0.05, off_s=(1, 2))
    run_sparsedata_case(rb, cuda, "f64", "L", "R", "T", "T", 64, 300, 700, 1.0, 2.0, fmt, 0.05, ma="S")
    monkeypatch.delenv("RBX_SPD_SLAB", raising=False)
    rng = np.random.default_rng(5)
    A = to_fmt(rb, cuda, random_sparse(rng, 2000, 1500, 0.02, np.float64), fmt)
    S = rb.DenseSkOp(rb.DenseDist(96, 2000), 3, dtype=torch.float64)
    outs = []
    for _ in range(3):
        B = torch.zeros(96 * 1500, dtype=torch.float64, device=cuda)
        rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 96, 1500, 2000, 1.0, S, 0, 0, A, 0, 0, 0.0, B, 96)
        outs.append(B.cpu().numpy())
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sparse_conversions_on_device(rb, cuda, dt):
    """conversions.hh:43-159 on the device (rbx_sparse_regroup_*): COO -> CSC / CSR from shuffled triplets, CSR <-> CSC,
    the zero-copy transpose views, empty rows / columns, an empty matrix; results equal scipy's canonical forms exactly."""
    import scipy.sparse as sp
    rng = np.random.default_rng(21)
    npdt = np.float32 if dt == "f32" else np.float64
    for (r, c, dens) in [(40, 70, 0.1), (300, 50, 0.3), (64, 64, 0.0), (5, 2000, 0.02)]:
        M = random_sparse(rng, r, c, dens, npdt)
        coo = sp.coo_matrix(M)
        perm = rng.permutation(coo.nnz)
        t = lambda a, d=None: torch.from_numpy(np.ascontiguousarray(a if d is None else a.astype(d))).to(cuda)
        A = rb.COOMatrix(r, c, t(coo.data[perm]), t(coo.row[perm], np.int64), t(coo.col[perm], np.int64))
        want_csc, want_csr = sp.csc_matrix(M), sp.csr_matrix(M)
        want_csc.sort_indices(); want_csr.sort_indices()
        for got, want, ptr, idx in [(rb.coo_to_csc(A), want_csc, "colptr", "rowidxs"), (rb.coo_to_csr(A), want_csr, "rowptr", "colidxs"),
                                    (rb.csr_to_csc(to_fmt(rb, cuda, M, "R")), want_csc, "colptr", "rowidxs"),
                                    (rb.csc_to_csr(to_fmt(rb, cuda, M, "C")), want_csr, "rowptr", "colidxs")]:
            assert np.array_equal(getattr(got, ptr).cpu().numpy(), want.indptr)
            assert np.array_equal(getattr(got, idx).cpu().numpy(), want.indices)
            assert np.array_equal(got.vals.cpu().numpy(), want.data)
        At = rb.transpose_as_csr(rb.coo_to_csc(A))
        assert (At.n_rows, At.n_cols) == (c, r) and np.array_equal(At.rowptr.cpu().numpy(), want_csc.indptr)


# ---------------------------------------------------------------------------------------------- wrappers
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sketch_vector(rb, cuda, dt):
    """test_sketch_vector.cc: y = alpha op(S) x + beta y with strides"""
    rng = np.random.default_rng(2)
    npdt = np.float32 if dt == "f32" else np.float64
    for (d, m, opS, incx, incy, off) in [(10, 30, "N", 1, 1, (0, 0)), (10, 30, "T", 2, 3, (0, 0)), (8, 25, "N", 3, 2, (2, 5)), (8, 25, "T", 1, 1, (1, 1))]:
        for mk in ("dense", "sparse"):
            Dr, Dc = d + off[0], m + off[1]
            lx, ly = (m, d) if opS == "N" else (d, m)
            x = rng.standard_normal(lx * incx).astype(npdt)
            y0 = rng.standard_normal(ly * incy).astype(npdt)
            if mk == "dense":
                S = rb.DenseSkOp(rb.DenseDist(Dr, Dc), 5, dtype=TDT[dt])
                Sfull = dense_S(dt, Dr, Dc, "G", "L", [0, 0, 0, 0], [5, 0], Dr, Dc, 0, 0)
            else:
                S = rb.SparseSkOp(rb.SparseDist(Dr, Dc, 3), 5, dtype=TDT[dt])
                Sfull = sparse_S_dense(dt, Dr, Dc, 3, "S", [0, 0, 0, 0], [5, 0])
            dx, dy = torch.from_numpy(x).to(cuda), torch.from_numpy(y0.copy()).to(cuda)
            rb.sketch_vector(rb.Op(opS), d, m, 1.5, S, off[0], off[1], dx, incx, -0.5, dy, incy)
            Sx = Sfull[off[0]:off[0] + d, off[1]:off[1] + m]
            Sop = Sx if opS == "N" else Sx.T
            got = dy.cpu().numpy()[::incy][:ly].astype(np.float64).reshape(-1, 1)
            check_bound(dt, got, y0[::incy][:ly].astype(np.float64).reshape(-1, 1), Sop, x[::incx][:lx].astype(np.float64).reshape(-1, 1),
                        1.5, -0.5, "L", ("vector", mk, opS))


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sketch_vector_gemv_path(rb, cuda, dt):
    """sketch_vector at sizes that take the generator-plus-dot-product kernels (rbx_sketch_dense_gemv.cu): both operator
    orientations (opS, major axis), unaligned offsets, strides, both families; 2 launches (partials + ordered reduce)."""
    rng = np.random.default_rng(12)
    npdt = np.float32 if dt == "f32" else np.float64
    for (d, m, opS, incx, incy, off, fam, ma) in [(300, 5000, "N", 1, 1, (0, 0), "G", "L"), (300, 5000, "N", 2, 3, (3, 5), "U", "L"),
                                                  (290, 3000, "T", 1, 2, (2, 7), "G", "L"), (300, 2500, "N", 1, 1, (1, 2), "G", "S"),
                                                  (130, 4099, "T", 3, 1, (0, 1), "U", "S")]:
        Dr, Dc = d + off[0] + 1, m + off[1] + 2
        lx, ly = (m, d) if opS == "N" else (d, m)
        x = rng.standard_normal(lx * incx).astype(npdt)
        y0 = rng.standard_normal(ly * incy).astype(npdt)
        S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), 5, dtype=TDT[dt])
        Sx = dense_S(dt, Dr, Dc, fam, ma, [0, 0, 0, 0], [5, 0], d, m, off[0], off[1])
        dx, dy = torch.from_numpy(x).to(cuda), torch.from_numpy(y0.copy()).to(cuda)
        l0 = rb.lib.rbx_launch_count()
        rb.sketch_vector(rb.Op(opS), d, m, 1.5, S, off[0], off[1], dx, incx, -0.5, dy, incy)
        assert rb.lib.rbx_launch_count() - l0 == 2
        Sop = Sx if opS == "N" else Sx.T
        got_flat = dy.cpu().numpy()
        got = got_flat[::incy][:ly].astype(np.float64).reshape(-1, 1)
        check_bound(dt, got, y0[::incy][:ly].astype(np.float64).reshape(-1, 1), Sop, x[::incx][:lx].astype(np.float64).reshape(-1, 1),
                    1.5, -0.5, "L", ("vector-gemv", opS, fam, ma))
        if incy > 1:      # the gaps of a strided y are untouched
            mask = np.ones(len(y0), bool); mask[::incy] = False
            assert np.array_equal(got_flat[mask], y0[mask])


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sketch_symmetric(rb, cuda, dt):
    """test_sketch_symmetric.cc:89-155: both sides, both layouts, both major axes; asymmetric input is rejected"""
    rng = np.random.default_rng(4)
    npdt = np.float32 if dt == "f32" else np.float64
    n, d = 23, 6
    for layout in "CR":
        for ma in "LS":
            M = rng.standard_normal((n, n))
            M = ((M + M.T) / 2).astype(npdt)
            lda = n + 2
            A = np.zeros(span(layout, n, n, lda), npdt)
            mat(layout, n, n, lda, A)[:] = M
            dA = torch.from_numpy(A).to(cuda)
            # right: B (n x d) = A S,  S is n x d
            S = rb.DenseSkOp(rb.DenseDist(n, d, "G", ma), 8, dtype=TDT[dt])
            ldb = n if layout == "C" else d
            dB = torch.zeros(n * d, dtype=TDT[dt], device=cuda)
            rb.sketch_symmetric(rb.Layout(layout), n, d, 1.0, dA, lda, S, 0, 0, 0.0, dB, ldb)
            Sx = dense_S(dt, n, d, "G", ma, [0, 0, 0, 0], [8, 0], n, d, 0, 0)
            check_bound(dt, mat(layout, n, d, ldb, dB.cpu().numpy()).astype(np.float64), np.zeros((n, d)), Sx, M.astype(np.float64), 1.0, 0.0, "R", "symm-right")
            # left: B (d x n) = S A,  S is d x n
            S = rb.DenseSkOp(rb.DenseDist(d, n, "G", ma), 8, dtype=TDT[dt])
            ldb = d if layout == "C" else n
            dB = torch.zeros(n * d, dtype=TDT[dt], device=cuda)
            rb.sketch_symmetric(rb.Layout(layout), d, n, 1.0, S, 0, 0, dA, lda, 0.0, dB, ldb)
            Sx = dense_S(dt, d, n, "G", ma, [0, 0, 0, 0], [8, 0], d, n, 0, 0)
            check_bound(dt, mat(layout, d, n, ldb, dB.cpu().numpy()).astype(np.float64), np.zeros((d, n)), Sx, M.astype(np.float64), 1.0, 0.0, "L", "symm-left")
    bad = torch.from_numpy(rng.standard_normal(n * n).astype(npdt)).to(cuda)
    with pytest.raises(rb.Error):
        rb.sketch_symmetric(rb.Layout.ColMajor, n, d, 1.0, bad, n, rb.DenseSkOp(rb.DenseDist(n, d), 8, dtype=TDT[dt]), 0, 0, 0.0,
                            torch.zeros(n * d, dtype=TDT[dt], device=cuda), n)


def test_host_pointer_entry_points(rb, cuda):
    """the rbx_host_* variants (host buffers, staged through HBM) give the same bits as the device-pointer calls"""
    rng = np.random.default_rng(6)
    A = torch.from_numpy(rng.standard_normal(300 * 20))
    S = rb.SparseSkOp(rb.SparseDist(16, 300, 4), 3)
    Bh = torch.zeros(16 * 20, dtype=torch.float64)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 16, 20, 300, 1.0, S, 0, 0, A, 300, 0.0, Bh, 16)
    Bd = torch.zeros(16 * 20, dtype=torch.float64, device=cuda)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 16, 20, 300, 1.0, S, 0, 0, A.to(cuda), 300, 0.0, Bd, 16)
    assert torch.allclose(Bh, Bd.cpu(), rtol=1e-13, atol=1e-13)
    Sd = rb.DenseSkOp(rb.DenseDist(16, 300), 3, dtype=torch.float64)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 16, 20, 300, 1.0, Sd, 0, 0, A, 300, 0.0, Bh, 16)
    rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 16, 20, 300, 1.0, Sd, 0, 0, A.to(cuda), 300, 0.0, Bd, 16)
    assert torch.equal(Bh, Bd.cpu())
    buf = torch.empty(16 * 300, dtype=torch.float32)
    rb.fill_dense(rb.DenseDist(16, 300), buf, rb.RNGState(3))
    rc, exp, _ = ol.fill_dense("f32", "R", 16, 300, "G", "L", 16, 300, 0, 0, [0, 0, 0, 0], [3, 0])
    assert np.array_equal(buf.numpy().view(np.uint32), exp.view(np.uint32))


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_host_pointer_padded_B_is_untouched_outside_the_view(rb, cuda, dt):
    """B as a window of a larger host array (ldb beyond the natural dimension, incy > 1) with beta == 0: the reference never
    touches the elements outside the rows x cols view, so they must come back bit-identical (the staged copy used
    to return uninitialised device memory there). Long and short unit-stride runs take different copy-back strategies."""
    rng = np.random.default_rng(77)
    npdt = np.float32 if dt == "f32" else np.float64
    for (layout, d, n, m, pad, sparse) in [("C", 100, 7, 300, 5, False), ("C", 8, 9, 200, 3, False), ("R", 6, 120, 200, 11, False),
                                           ("R", 40, 5, 300, 2, True), ("C", 128, 6, 400, 64, True)]:
        ldb = (d if layout == "C" else n) + pad
        lda = m if layout == "C" else n
        A = torch.from_numpy(rng.standard_normal(m * n).astype(npdt))
        B0 = rng.standard_normal(span(layout, d, n, ldb)).astype(npdt)
        B0[::3] = np.nan                                       # NaN in B with beta == 0 must not propagate either (util.hh:51-59)
        Bh = torch.from_numpy(B0.copy())
        S = rb.SparseSkOp(rb.SparseDist(d, m, 4), 3) if sparse else rb.DenseSkOp(rb.DenseDist(d, m), 3, dtype=TDT[dt])
        rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, lda, 0.0, Bh, ldb)
        Bd = torch.from_numpy(B0.copy()).to(cuda)
        rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A.to(cuda), lda, 0.0, Bd, ldb)
        got, dev = Bh.numpy(), Bd.cpu().numpy()
        mask = np.ones(len(B0), bool)
        mask[mat(layout, d, n, ldb, np.arange(len(B0))).ravel()] = False
        bits = np.uint32 if dt == "f32" else np.uint64
        assert np.array_equal(got[mask].view(bits), B0[mask].view(bits)), (layout, d, n, pad)
        assert np.isfinite(got[~mask]).all()
        assert np.allclose(got[~mask], dev[~mask], rtol=1e-5 if dt == "f32" else 1e-12, atol=1e-5 if dt == "f32" else 1e-12)
    # sketch_vector with a strided y on the host
    d, m, incy = 300, 5000, 3
    x = torch.from_numpy(rng.standard_normal(m).astype(npdt))
    y0 = rng.standard_normal(d * incy).astype(npdt)
    y = torch.from_numpy(y0.copy())
    rb.sketch_vector(rb.Op.NoTrans, d, m, 1.0, rb.DenseSkOp(rb.DenseDist(d, m), 5, dtype=TDT[dt]), 0, 0, x, 1, 0.0, y, incy)
    gap = np.ones(len(y0), bool); gap[::incy] = False
    assert np.array_equal(y.numpy()[gap], y0[gap]) and not np.array_equal(y.numpy()[::incy], y0[::incy])


@pytest.mark.parametrize("fmt", ["O", "R", "C"])
@pytest.mark.parametrize("back", ["whole", "blocks"])
def test_host_pointer_sketch_sparse(rb, cuda, fmt, back, monkeypatch):
    """rbx_host_sketch_sparsedata_*: the sparse matrix and B in host memory give the device call's bits; B as a window of a
    larger host array (ldb beyond the natural dimension) keeps its padding, with beta == 0 (NaNs in B must not propagate)
    and with beta != 0; argument errors are reported without touching B. back = "blocks": the result returns in eight column
    blocks on the auxiliary stream while later blocks are gathered (what results of 256 MB and more do; RBX_SPD_BACK_MB = 0
    forces it here), contiguous and strided (padded) views."""
    if back == "blocks":
        monkeypatch.setenv("RBX_SPD_BACK_MB", "0")
    rng = np.random.default_rng(91)
    for dt in ("f32", "f64"):
        npdt = np.float32 if dt == "f32" else np.float64
        bits = np.uint32 if dt == "f32" else np.uint64
        for (side, layout, d, n, m, pad, beta) in [("L", "C", 24, 300, 500, 0, 0.0), ("L", "C", 24, 40, 500, 7, 0.0), ("L", "R", 12, 130, 260, 5, -0.5),
                                                  ("R", "C", 16, 200, 90, 3, 0.0), ("R", "R", 16, 200, 90, 0, 1.5), ("L", "C", 72, 64, 300, 5, 0.0),
                                                  ("R", "R", 80, 37, 120, 9, -1.0), ("L", "C", 8, 5, 60, 0, 0.0)]:
            # left: B (d x n) = S (d x m) A (m x n);  right: B (m' x d) = A (m' x n') S (n' x d) with m' = n, n' = m here
            rowsA, colsA = (m, n) if side == "L" else (n, m)
            rowsB, colsB = (d, n) if side == "L" else (n, d)
            Afull = random_sparse(rng, rowsA, colsA, 0.05, npdt)
            Ad = to_fmt(rb, cuda, Afull, fmt)
            i0, i1 = Ad._idx()
            Ah = type(Ad)(Ad.n_rows, Ad.n_cols, Ad.vals.cpu(), i0.cpu(), i1.cpu())
            ldb = (rowsB if layout == "C" else colsB) + pad
            B0 = rng.standard_normal(span(layout, rowsB, colsB, ldb)).astype(npdt)
            if beta == 0.0:
                B0[::3] = np.nan
            S = rb.DenseSkOp(rb.DenseDist(d, m) if side == "L" else rb.DenseDist(m, d), 13, dtype=TDT[dt])
            L, O = rb.Layout(layout), rb.Op
            Bh, Bd = torch.from_numpy(B0.copy()), torch.from_numpy(B0.copy()).to(cuda)
            for (A_, B_) in ((Ah, Bh), (Ad, Bd)):
                if side == "L":
                    rb.sketch_sparse(L, O.NoTrans, O.NoTrans, d, n, m, 2.0, S, 0, 0, A_, 0, 0, beta, B_, ldb)
                else:
                    rb.sketch_sparse(L, O.NoTrans, O.NoTrans, n, d, m, 2.0, A_, 0, 0, S, 0, 0, beta, B_, ldb)
            got, dev = Bh.numpy(), Bd.cpu().numpy()
            assert np.array_equal(got.view(bits), dev.view(bits)), (dt, side, layout, fmt, pad)
            mask = np.ones(len(B0), bool)
            mask[mat(layout, rowsB, colsB, ldb, np.arange(len(B0))).ravel()] = False
            assert np.array_equal(got[mask].view(bits), B0[mask].view(bits)) and np.isfinite(got[~mask]).all()
    # the reference's dimension check (sksp.hh:175-178) comes back as Error before anything is written
    Bh = torch.full((24 * 300,), 7.0, dtype=torch.float64)
    with pytest.raises(rb.Error):
        rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, 24, 300, 501, 1.0, rb.DenseSkOp(rb.DenseDist(24, 501), 1, dtype=torch.float64),
                         0, 0, Ah, 0, 0, 0.0, Bh, 24)
    assert float(Bh.min()) == 7.0 and float(Bh.max()) == 7.0


@pytest.mark.parametrize("layout", ["C", "R"])
def test_host_pointer_pipelined_slabs(rb, cuda, layout):
    """host matrices of >= 1 GiB are sketched slab by slab (copy of slab i+1 overlaps the sketch of slab i, co_s advanced per
    slab, beta carried by the first slab): same result as the device-resident call up to the summation order"""
    d, n, m = 256, 1024, (1 << 17) + 1000            # 1 GiB + a ragged last slab
    A = torch.empty(m * n, dtype=torch.float64, device=cuda)
    rb.fill_dense(rb.Layout(layout), rb.DenseDist(m, n), m, n, 0, 0, A, rb.RNGState(99))
    lda, ldb = (m, d) if layout == "C" else (n, n)
    S = rb.SparseSkOp(rb.SparseDist(d, m + 24, 8), 1997)
    B0 = torch.ones(d * n, dtype=torch.float64)
    Bd = B0.to(cuda)
    rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, S, 0, 24, A, lda, -0.5, Bd, ldb)
    Bh = B0.clone()
    rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, S, 0, 24, A.cpu(), lda, -0.5, Bh, ldb)
    scale = float(Bd.abs().max())
    assert float((Bh - Bd.cpu()).abs().max()) <= 1e-12 * scale
    # the same pipeline serves dense operators (co_s advanced per slab selects the operator's columns of that slab)
    Sd = rb.DenseSkOp(rb.DenseDist(d, m + 24), 5, dtype=torch.float64)
    Bd = B0.to(cuda)
    rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, Sd, 0, 24, A, lda, -0.5, Bd, ldb)
    Bh = B0.clone()
    rb.sketch_general(rb.Layout(layout), rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.5, Sd, 0, 24, A.cpu(), lda, -0.5, Bh, ldb)
    scale = float(Bd.abs().max())
    assert float((Bh - Bd.cpu()).abs().max()) <= 1e-11 * scale


def test_implicit_dense_sketch_allocates_nothing_proportional_to_the_operator(rb, cuda):
    """north-star item 2 / VERDICT r1 #4: an implicit DenseSkOp is never materialised. The private workspace pool's high-water
    mark (rbx_workspace_stats) after sketches whose operators would take 4.3 GB (float) and 2.1 GB (double) stays under the
    fixed panel budget: two k-panels of at most 512 MB (float, tcgen05) / 256 MB (double, DMMA) plus the split-K partials."""
    import ctypes
    def high_water(reset):
        un, uh, rn = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        assert rb.lib.rbx_workspace_stats(ctypes.byref(un), ctypes.byref(uh), ctypes.byref(rn), 1 if reset else 0) == 0
        return uh.value
    for tdt, d, n, m, fam, budget in ((torch.float32, 1024, 256, 1 << 20, "G", 1100 << 20), (torch.float32, 1024, 256, 1 << 20, "U", 64 << 20),
                                     (torch.float64, 512, 256, 1 << 19, "G", 600 << 20)):
        A = torch.empty(m * n, dtype=tdt, device=cuda)
        rb.fill_dense(rb.DenseDist(m, n, "U"), A, rb.RNGState(4))
        B = torch.zeros(d * n, dtype=tdt, device=cuda)
        S = rb.DenseSkOp(rb.DenseDist(d, m, fam), 9, dtype=tdt)
        torch.cuda.synchronize()
        high_water(True)
        rb.sketch_general(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, m, 0.0, B, d)
        torch.cuda.synchronize()
        hw = high_water(True)
        operator_bytes = d * m * A.element_size()
        assert hw <= budget and hw < operator_bytes // 3, (fam, tdt, hw, operator_bytes)
        assert torch.isfinite(B).all() and float(B.abs().max()) > 0
        del A, B
