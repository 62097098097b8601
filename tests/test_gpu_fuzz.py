"""Randomized differential tests on the GPU: seeded random shapes, layouts, transposes, operator windows, leading dimensions,
scalars, families and major axes through the public calls, checked against the oracle (fills: bit for bit; sketches: the
reference's componentwise bound, test/test_matmul_cores/linop_common.hh:257-263). The hand-picked cases elsewhere in tests/
sit on the kernels' envelopes; these draws look for what nobody thought of. Every dense draw runs twice: with the shipping
dispatch (small shapes take the SIMT kernel) and with RBX_DENSE_MIN_MACS=0, which sends every shape the tensor-core kernels
accept to them (tcgen05 for float, DMMA for double, their k-major copy / split-K / k-panel routes included).
The seeds are fixed: a failure names its draw and reproduces. RBX_FUZZ_BLOCKS=k runs k times as many blocks of draws."""
import os

import numpy as np
import pytest
import torch

import oracle_lib as ol
from test_gpu_fill import bits
from test_gpu_sketch import (TDT, check_bound, dense_S, run_dense_case, run_sparsedata_case, run_sparseop_case, sparse_S_dense)

pytestmark = pytest.mark.gpu

MORE = int(os.environ.get("RBX_FUZZ_BLOCKS", "1"))
SCALARS = [(1.0, 0.0), (1.0, 0.0), (1.5, -0.5), (-0.7, 1.0), (0.0, 2.0), (2.0, 0.25)]


def logint(rng, lo, hi):
    """integer in [lo, hi], log-uniform: small and large sizes are equally likely"""
    return int(min(hi, max(lo, round(float(np.exp(rng.uniform(np.log(lo), np.log(hi + 0.49))))))))


def pick(rng, seq):
    return seq[int(rng.integers(len(seq)))]


def small_offsets(rng):
    return (int(rng.integers(0, 9)) if rng.random() < 0.6 else 0, int(rng.integers(0, 9)) if rng.random() < 0.6 else 0)


# ---------------------------------------------------------------------------------------------- dense operators
@pytest.mark.parametrize("macs", ["default", "0"])
@pytest.mark.parametrize("block", range(6 * MORE))
def test_fuzz_sketch_general_dense(rb, cuda, monkeypatch, block, macs):
    if macs == "0":
        monkeypatch.setenv("RBX_DENSE_MIN_MACS", "0")
    rng = np.random.default_rng(1000 + block)
    multi = 0
    for it in range(14):
        dt, side, layout = pick(rng, ["f32", "f64"]), pick(rng, "LR"), pick(rng, "CR")
        opS, opA = pick(rng, "NNT"), pick(rng, "NNT")
        d, n, m = logint(rng, 1, 300), logint(rng, 1, 300), logint(rng, 1 if rng.random() < 0.3 else 16, 6000)
        if rng.random() < 0.2:
            d, m = m % 700 + 1, d            # lifts: the operator's long axis is not the contraction
        if side == "R":
            n, m = m, n                      # B (m x d) = op(A) (m x n) op(S) (n x d): n is contracted there
        alpha, beta = pick(rng, SCALARS)
        draw = dict(dt=dt, side=side, layout=layout, opS=opS, opA=opA, d=d, n=n, m=m, alpha=alpha, beta=beta, off=small_offsets(rng),
                    D_extra=small_offsets(rng), fam=pick(rng, "GU"), ma=pick(rng, "LS"), lda_pad=pick(rng, [0, 0, 1, 2, 3, 4, 8]),
                    ldb_pad=pick(rng, [0, 0, 1, 3, 4]), key=(int(rng.integers(0, 2 ** 32)), int(rng.integers(0, 2 ** 32))),
                    prefilled=bool(rng.random() < 0.2), host=bool(rng.random() < 0.2))
        if draw["host"]:
            draw["prefilled"] = False        # host-pointer sketches take implicit operators
        l0 = rb.lib.rbx_launch_count()
        try:
            run_dense_case(rb, cuda, **draw)
        except Exception as e:      # name the draw
            raise AssertionError(f"dense draw {block}.{it} (RBX_DENSE_MIN_MACS={macs}): {draw}: {e}") from e
        multi += (rb.lib.rbx_launch_count() - l0 - (1 if draw["prefilled"] else 0)) >= 2
    if macs == "0":
        assert multi >= 2, f"RBX_DENSE_MIN_MACS=0 sent only {multi} of 14 draws to the tensor-core kernels"


# ---------------------------------------------------------------------------------------------- sparse operators
@pytest.mark.parametrize("block", range(6 * MORE))
def test_fuzz_sketch_general_sparse_operator(rb, cuda, block):
    rng = np.random.default_rng(2000 + block)
    for it in range(12):
        dt, side, layout = pick(rng, ["f32", "f64"]), pick(rng, "LR"), pick(rng, "CR")
        opS, opA = pick(rng, "NNT"), pick(rng, "NNT")
        d, n, m = logint(rng, 1, 2500), logint(rng, 1, 160), logint(rng, 1, 20000)
        while d * m > 1.5e7:               # the checker builds the operator densely
            d, m = max(1, d // 2), max(1, m // 2)
        if rng.random() < 0.2:
            d, m = m, d                     # lift
        off, extra = small_offsets(rng), small_offsets(rng)
        if rng.random() < 0.3:
            off = (off[0], int(rng.integers(0, 3000)))
        alpha, beta = pick(rng, SCALARS)
        if side == "R":                     # B (m x d) = op(A) (m x n) op(S) (n x d): n is the long axis there
            n, m = m, n
            while n * d > 1.5e7:
                n = max(1, n // 2)
        rowsS, colsS = ((d, m) if opS == "N" else (m, d)) if side == "L" else ((n, d) if opS == "N" else (d, n))
        short = min(rowsS + off[0] + extra[0], colsS + off[1] + extra[1])
        k = min(short, pick(rng, [1, 2, 3, 4, 7, 8, 8, 8, 13, 32, 33, 40]))
        draw = dict(dt=dt, side=side, layout=layout, opS=opS, opA=opA, d=d, n=n, m=m, alpha=alpha, beta=beta, k=k, ma=pick(rng, "SSL"),
                    D_extra=extra, off=off, key=int(rng.integers(0, 2 ** 31)), lda_pad=pick(rng, [0, 0, 1, 2, 4]),
                    filled=bool(rng.random() < 0.25), host=bool(rng.random() < 0.2))
        try:
            run_sparseop_case(rb, cuda, **draw)
        except Exception as e:
            raise AssertionError(f"sparse-operator draw {block}.{it}: {draw}: {e}") from e


# ---------------------------------------------------------------------------------------------- sparse data
@pytest.mark.parametrize("block", range(4 * MORE))
def test_fuzz_sketch_sparse(rb, cuda, block):
    rng = np.random.default_rng(3000 + block)
    for it in range(12):
        dt, side, layout, fmt = pick(rng, ["f32", "f64"]), pick(rng, "LR"), pick(rng, "CR"), pick(rng, "ORC")
        d, n, m = logint(rng, 1, 700), logint(rng, 1, 500), logint(rng, 1, 900)
        alpha, beta = pick(rng, SCALARS)
        draw = dict(dt=dt, side=side, layout=layout, opS=pick(rng, "NNT"), opA=pick(rng, "NNT"), d=d, n=n, m=m, alpha=alpha, beta=beta,
                    fmt=fmt, density=pick(rng, [0.0, 0.004, 0.02, 0.1, 0.5]), off_s=small_offsets(rng),
                    off_a=small_offsets(rng) if fmt == "O" else (0, 0), fam=pick(rng, "GU"), ma=pick(rng, "LS"), host=bool(rng.random() < 0.25))
        try:
            run_sparsedata_case(rb, cuda, **draw)
        except Exception as e:
            raise AssertionError(f"sparse-data draw {block}.{it}: {draw}: {e}") from e


# ---------------------------------------------------------------------------------------------- vectors
@pytest.mark.parametrize("block", range(4 * MORE))
def test_fuzz_sketch_vector(rb, cuda, block):
    rng = np.random.default_rng(4000 + block)
    for it in range(14):
        dt, opS, mk = pick(rng, ["f32", "f64"]), pick(rng, "NT"), pick(rng, ["dense", "dense", "sparse"])
        npdt = np.float32 if dt == "f32" else np.float64
        d, m = logint(rng, 1, 400), logint(rng, 1, 30000)
        if rng.random() < 0.3:
            d, m = m % 5000 + 1, d
        while d * m > 4e6:
            d, m = max(1, d // 2), max(1, m // 2)
        off, extra = small_offsets(rng), small_offsets(rng)
        incx, incy = pick(rng, [1, 1, 2, 3]), pick(rng, [1, 1, 2, 5])
        alpha, beta = pick(rng, SCALARS)
        Dr, Dc = d + off[0] + extra[0], m + off[1] + extra[1]
        fam, ma, key = pick(rng, "GU"), pick(rng, "LS"), int(rng.integers(0, 2 ** 31))
        lx, ly = (m, d) if opS == "N" else (d, m)
        x = rng.standard_normal(lx * incx).astype(npdt)
        y0 = rng.standard_normal(ly * incy).astype(npdt)
        if mk == "dense":
            S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), key, dtype=TDT[dt])
            Sx = dense_S(dt, Dr, Dc, fam, ma, [0, 0, 0, 0], [key, 0], d, m, off[0], off[1])
        else:
            k = min(Dr, Dc, pick(rng, [1, 3, 8, 8, 33]))
            S = rb.SparseSkOp(rb.SparseDist(Dr, Dc, k, ma), key, dtype=TDT[dt])
            Sx = sparse_S_dense(dt, Dr, Dc, k, ma, [0, 0, 0, 0], [key, 0])[off[0]:off[0] + d, off[1]:off[1] + m]
        what = dict(dt=dt, opS=opS, kind=mk, d=d, m=m, off=off, extra=extra, incx=incx, incy=incy, fam=fam, ma=ma, key=key, alpha=alpha, beta=beta)
        try:
            dx, dy = torch.from_numpy(x).to(cuda), torch.from_numpy(y0.copy()).to(cuda)
            rb.sketch_vector(rb.Op(opS), d, m, alpha, S, off[0], off[1], dx, incx, beta, dy, incy)
            got_flat = dy.cpu().numpy()
            Sop = Sx if opS == "N" else Sx.T
            check_bound(dt, got_flat[::incy][:ly].astype(np.float64).reshape(-1, 1), y0[::incy][:ly].astype(np.float64).reshape(-1, 1), Sop,
                        x[::incx][:lx].astype(np.float64).reshape(-1, 1), alpha, beta, "L", "vector")
            if incy > 1:
                mask = np.ones(len(y0), bool); mask[::incy] = False
                assert np.array_equal(got_flat[mask], y0[mask]), "gaps of a strided y were written"
        except Exception as e:
            raise AssertionError(f"vector draw {block}.{it}: {what}: {e}") from e


# ---------------------------------------------------------------------------------------------- fills
def random_counter(rng):
    """mostly small counters; sometimes one whose low words are about to carry"""
    r = rng.random()
    if r < 0.5:
        return [int(rng.integers(0, 1000)), 0, 0, 0]
    if r < 0.8:
        return [0xFFFFFFFF - int(rng.integers(0, 5000)), 0xFFFFFFFF if rng.random() < 0.5 else int(rng.integers(0, 2 ** 32)), int(rng.integers(0, 2 ** 32)), 0]
    return [int(rng.integers(0, 2 ** 32)) for _ in range(3)] + [int(rng.integers(0, 2 ** 31))]


@pytest.mark.parametrize("block", range(4 * MORE))
def test_fuzz_fill_dense_bits(rb, cuda, block):
    rng = np.random.default_rng(5000 + block)
    for it in range(20):
        dt, layout, fam, ma = pick(rng, ["f32", "f64"]), pick(rng, "CR"), pick(rng, "GU"), pick(rng, "LS")
        Dr, Dc = logint(rng, 1, 3000), logint(rng, 1, 3000)
        while Dr * Dc > 3e6:
            Dr, Dc = max(1, Dr // 2), max(1, Dc // 2)
        if rng.random() < 0.4:
            rows, cols, ro, co = Dr, Dc, 0, 0
        else:
            rows, cols = int(rng.integers(1, Dr + 1)), int(rng.integers(1, Dc + 1))
            ro, co = int(rng.integers(0, Dr - rows + 1)), int(rng.integers(0, Dc - cols + 1))
        ctr, key = random_counter(rng), [int(rng.integers(0, 2 ** 32)), int(rng.integers(0, 2 ** 32))]
        what = dict(dt=dt, layout=layout, fam=fam, ma=ma, D=(Dr, Dc), window=(rows, cols, ro, co), ctr=ctr, key=key)
        buf = torch.full((rows * cols,), float("nan"), dtype=TDT[dt], device=cuda)
        nxt = rb.fill_dense(rb.Layout(layout), rb.DenseDist(Dr, Dc, fam, ma), rows, cols, ro, co, buf, rb.RNGState(key, ctr))
        rc, exp, exp_next = ol.fill_dense(dt, layout, Dr, Dc, fam, ma, rows, cols, ro, co, ctr, key)
        assert rc == 0
        assert np.array_equal(bits(buf.cpu().numpy()), bits(exp)), f"fill_dense draw {block}.{it}: {what}"
        assert list(nxt.counter) == list(exp_next), f"fill_dense next state, draw {block}.{it}: {what}"


@pytest.mark.parametrize("block", range(3 * MORE))
def test_fuzz_fill_sparse_bits(rb, cuda, block):
    rng = np.random.default_rng(6000 + block)
    for it in range(16):
        dt, ma = pick(rng, ["f32", "f64"]), pick(rng, "SSL")
        r, c = logint(rng, 1, 4000), logint(rng, 1, 60000)
        if rng.random() < 0.4:
            r, c = c, r
        k = min(r, c, pick(rng, [1, 2, 3, 8, 8, 17, 32, 33, 100, 700]))
        while k * (max(r, c) if ma == "S" else min(r, c)) > 3e6:
            k = max(1, k // 2)
        key = int(rng.integers(0, 2 ** 31))
        what = dict(dt=dt, ma=ma, shape=(r, c), vec_nnz=k, key=key)
        S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), key, dtype=TDT[dt])
        rb.fill_sparse(S)
        rc, rows, cols, vals, exp_next = ol.fill_sparse(dt, "i64", r, c, k, ma, [0, 0, 0, 0], [key, 0])
        assert rc == 0
        assert np.array_equal(S.rows.cpu().numpy(), rows) and np.array_equal(S.cols.cpu().numpy(), cols), f"fill_sparse indices, draw {block}.{it}: {what}"
        assert np.array_equal(bits(S.vals.cpu().numpy()), bits(vals)), f"fill_sparse values, draw {block}.{it}: {what}"
        assert list(S.next_state.counter) == list(exp_next), f"fill_sparse next state, draw {block}.{it}: {what}"


# ---------------------------------------------------------------------------------------------- medium sizes, shipping dispatch
@pytest.mark.parametrize("block", range(2 * MORE))
def test_fuzz_medium_sizes(rb, cuda, block):
    """draws large enough (1e8 .. 1e10 multiply-adds) for the shipping dispatch to pick the tensor-core and streaming kernels"""
    rng = np.random.default_rng(7000 + block)
    for it in range(6):
        dt, side, layout = pick(rng, ["f32", "f64"]), pick(rng, "LR"), pick(rng, "CR")
        alpha, beta = pick(rng, SCALARS)
        d, n, m = logint(rng, 8, 700), logint(rng, 2, 900), logint(rng, 2000, 60000)
        while d * m > 2e7:
            d, m = max(1, d * 2 // 3), max(1, m * 2 // 3)
        if side == "R":
            n, m = m, n
        draw = dict(dt=dt, side=side, layout=layout, opS=pick(rng, "NNT"), opA=pick(rng, "NNT"), d=d, n=n, m=m, alpha=alpha, beta=beta,
                    off=small_offsets(rng), D_extra=small_offsets(rng), fam=pick(rng, "GU"), ma=pick(rng, "LS"), lda_pad=pick(rng, [0, 0, 1, 3, 4]),
                    ldb_pad=pick(rng, [0, 0, 1, 4]), key=(int(rng.integers(0, 2 ** 32)), 0), prefilled=bool(rng.random() < 0.25))
        try:
            run_dense_case(rb, cuda, **draw)
        except Exception as e:
            raise AssertionError(f"medium dense draw {block}.{it}: {draw}: {e}") from e
        k = pick(rng, [1, 4, 8, 8, 8, 16, 32, 40])
        sdraw = dict(dt=dt, side=side, layout=layout, opS=draw["opS"], opA=draw["opA"], d=d, n=n, m=m, alpha=alpha, beta=beta, k=min(k, d),
                     ma=pick(rng, "SSSL"), D_extra=draw["D_extra"], off=draw["off"], key=int(rng.integers(0, 2 ** 31)), lda_pad=draw["lda_pad"])
        try:
            run_sparseop_case(rb, cuda, **sdraw)
        except Exception as e:
            raise AssertionError(f"medium sparse-operator draw {block}.{it}: {sdraw}: {e}") from e


# ---------------------------------------------------------------------------------------------- spmm, symmetric
@pytest.mark.parametrize("block", range(3 * MORE))
def test_fuzz_left_right_spmm(rb, cuda, block):
    """sparse_data::left_spmm / right_spmm (spmm_dispatch.hh:48-200) with random formats, layouts, transposes, scalars, leading
    dimensions and (COO) sub-matrix offsets, against a dense product"""
    from test_gpu_sketch import mat, random_sparse, span, to_fmt
    rng = np.random.default_rng(8000 + block)
    for it in range(10):
        dt, fmt, layout, side = pick(rng, ["f32", "f64"]), pick(rng, "ORC"), pick(rng, "CR"), pick(rng, "LR")
        npdt = np.float32 if dt == "f32" else np.float64
        opA, opB = pick(rng, "NNT"), pick(rng, "NNT")         # opA: the sparse operand, opB: the dense one
        d, n, m = logint(rng, 1, 400), logint(rng, 1, 300), logint(rng, 1, 3000)
        alpha, beta = pick(rng, SCALARS)
        off = small_offsets(rng) if fmt == "O" else (0, 0)
        density = pick(rng, [0.0, 0.01, 0.05, 0.3])
        what = dict(dt=dt, fmt=fmt, layout=layout, side=side, opA=opA, opB=opB, d=d, n=n, m=m, alpha=alpha, beta=beta, off=off, density=density)
        try:
            # left: C (d x n) = op(A) (d x m) op(B) (m x n); right: C (n x d) = op(B) (n x m) op(A) (m x d)
            rowsA, colsA = ((d, m) if opA == "N" else (m, d)) if side == "L" else ((m, d) if opA == "N" else (d, m))
            rowsB, colsB = ((m, n) if opB == "N" else (n, m)) if side == "L" else ((n, m) if opB == "N" else (m, n))
            rowsC, colsC = (d, n) if side == "L" else (n, d)
            Afull = random_sparse(rng, rowsA + off[0] + 2, colsA + off[1] + 1, density, npdt) if fmt == "O" else random_sparse(rng, rowsA, colsA, density, npdt)
            A = to_fmt(rb, cuda, Afull, fmt)
            ldb = (rowsB if layout == "C" else colsB) + pick(rng, [0, 0, 1, 3])
            ldc = (rowsC if layout == "C" else colsC) + pick(rng, [0, 0, 2])
            Bflat = rng.standard_normal(span(layout, rowsB, colsB, ldb)).astype(npdt)
            C0 = rng.standard_normal(span(layout, rowsC, colsC, ldc)).astype(npdt)
            dB, dC = torch.from_numpy(Bflat).to(cuda), torch.from_numpy(C0.copy()).to(cuda)
            L, O = rb.Layout(layout), rb.Op
            if side == "L":
                rb.left_spmm(L, O(opA), O(opB), d, n, m, alpha, A, off[0], off[1], dB, ldb, beta, dC, ldc)
            else:
                rb.right_spmm(L, O(opB), O(opA), n, d, m, alpha, dB, ldb, A, off[0], off[1], beta, dC, ldc)
            Ax = Afull[off[0]:off[0] + rowsA, off[1]:off[1] + colsA].astype(np.float64)
            Aop = Ax if opA == "N" else Ax.T
            Bop = mat(layout, rowsB, colsB, ldb, Bflat).astype(np.float64)
            Bop = Bop if opB == "N" else Bop.T
            got_flat = dC.cpu().numpy()
            got = mat(layout, rowsC, colsC, ldc, got_flat).astype(np.float64)
            C0m = mat(layout, rowsC, colsC, ldc, C0).astype(np.float64)
            # the componentwise bound with the sparse operand in the operator's place
            check_bound(dt, got, C0m, Aop, Bop, alpha, beta, side, "spmm")
            mask = np.ones(len(C0), bool)
            mask[mat(layout, rowsC, colsC, ldc, np.arange(len(C0))).ravel()] = False
            assert np.array_equal(got_flat[mask], C0[mask]), "padding of C was written"
        except Exception as e:
            raise AssertionError(f"spmm draw {block}.{it}: {what}: {e}") from e


@pytest.mark.parametrize("block", range(2 * MORE))
def test_fuzz_sketch_symmetric(rb, cuda, block):
    """sketch_symmetric (sksy.hh) with random sizes, layouts, sides, families, major axes, operator windows and scalars"""
    from test_gpu_sketch import mat, span
    rng = np.random.default_rng(9000 + block)
    for it in range(10):
        dt, layout, side = pick(rng, ["f32", "f64"]), pick(rng, "CR"), pick(rng, "LR")
        npdt = np.float32 if dt == "f32" else np.float64
        n, d = logint(rng, 1, 700), logint(rng, 1, 200)
        fam, ma, key = pick(rng, "GU"), pick(rng, "LS"), int(rng.integers(0, 2 ** 31))
        alpha, beta = pick(rng, SCALARS)
        off, extra = small_offsets(rng), small_offsets(rng)
        what = dict(dt=dt, layout=layout, side=side, n=n, d=d, fam=fam, ma=ma, key=key, alpha=alpha, beta=beta, off=off, extra=extra)
        try:
            M = rng.standard_normal((n, n)); M = ((M + M.T) / 2).astype(npdt)
            lda = n + pick(rng, [0, 0, 1, 2])
            A = np.zeros(span(layout, n, n, lda), npdt)
            mat(layout, n, n, lda, A)[:] = M
            dA = torch.from_numpy(A).to(cuda)
            rowsS, colsS = (d, n) if side == "L" else (n, d)
            rowsB, colsB = (d, n) if side == "L" else (n, d)
            Dr, Dc = rowsS + off[0] + extra[0], colsS + off[1] + extra[1]
            S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), key, dtype=TDT[dt])
            ldb = (rowsB if layout == "C" else colsB) + pick(rng, [0, 0, 3])
            B0 = rng.standard_normal(span(layout, rowsB, colsB, ldb)).astype(npdt)
            dB = torch.from_numpy(B0.copy()).to(cuda)
            if side == "L":
                rb.sketch_symmetric(rb.Layout(layout), d, n, alpha, S, off[0], off[1], dA, lda, beta, dB, ldb)
            else:
                rb.sketch_symmetric(rb.Layout(layout), n, d, alpha, dA, lda, S, off[0], off[1], beta, dB, ldb)
            Sx = dense_S(dt, Dr, Dc, fam, ma, [0, 0, 0, 0], [key, 0], rowsS, colsS, off[0], off[1])
            got = mat(layout, rowsB, colsB, ldb, dB.cpu().numpy()).astype(np.float64)
            check_bound(dt, got, mat(layout, rowsB, colsB, ldb, B0).astype(np.float64), Sx, M.astype(np.float64), alpha, beta, side, "symmetric")
        except Exception as e:
            raise AssertionError(f"symmetric draw {block}.{it}: {what}: {e}") from e
