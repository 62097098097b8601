"""Parity links that the round-1 review found soft, closed on the GPU box itself:
  (i)   device sketches against the COMPILED REFERENCE's own lskge3 / rskge3 / lskges / rskges / lsksp3 (oracle/_ref,
        which="ref"), not only against the C restatement;
  (ii)  C3 at 2^32: 64 random 2^20-entry tiles of the device fill against the compiled reference's fill_dense of the same
        windows, bit for bit, plus every lane phase at the buffer's edges;
  (iii) the exhaustive 2^32 sweep that makes the Gaussian bit-exactness provable: every raw word through the
        device's sincosf / logf models against THIS host's libm (RBX_SWEEP_STRIDE > 1 subsamples; default = all 2^32);
  (iv)  the oracle pins (tests/test_oracle_pins.py: restatement <-> compiled reference <-> goldens) executed on the GPU
        box's libm as part of `-m gpu`.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EPS = {"f32": 2.0 ** -23, "f64": 2.0 ** -52}
TDT = {"f32": torch.float32, "f64": torch.float64}
needs_ref = pytest.mark.skipif(ol.REF is None, reason="compiled reference (oracle/_ref) not built")


@needs_ref
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_dense_sketches_against_the_compiled_reference(rb, cuda, dt):
    """test_lskge3.cc / test_rskge3.cc shapes: the device result vs the reference's own result on the same inputs"""
    rng = np.random.default_rng(5)
    npdt = np.float32 if dt == "f32" else np.float64
    cases = [("L", "C", "N", "N", 7, 5, 19, (0, 0)), ("L", "R", "T", "N", 19, 4, 7, (3, 1)), ("L", "C", "N", "T", 130, 70, 517, (2, 5)),
             ("R", "C", "N", "N", 90, 64, 400, (1, 1)), ("R", "R", "T", "T", 33, 17, 129, (0, 3)), ("L", "C", "N", "N", 128, 256, 4096, (0, 0))]
    for side, layout, opS, opA, p0, p1, p2, off in cases:
        if side == "L":
            d, n, m = p0, p1, p2
            rS, cS = (d, m) if opS == "N" else (m, d)
            rA, cA = (m, n) if opA == "N" else (n, m)
            rB, cB, K = d, n, m
        else:
            m, d, n = p0, p1, p2
            rS, cS = (n, d) if opS == "N" else (d, n)
            rA, cA = (m, n) if opA == "N" else (n, m)
            rB, cB, K = m, d, n
        Dr, Dc = rS + off[0] + 2, cS + off[1] + 3
        lda, ldb = (rA if layout == "C" else cA) + 1, (rB if layout == "C" else cB) + 2
        nA = (cA - 1) * lda + rA if layout == "C" else (rA - 1) * lda + cA
        nB = (cB - 1) * ldb + rB if layout == "C" else (rB - 1) * ldb + cB
        A = rng.standard_normal(nA).astype(npdt)
        B0 = rng.standard_normal(nB).astype(npdt)
        for fam, ma in (("G", "L"), ("U", "S")):
            want = B0.copy()
            assert ol.sketch_dense(dt, side, layout, opS, opA, p0, p1, p2, 1.5, Dr, Dc, fam, ma, [0, 0, 0, 0], [3, 0], off[0], off[1], A, lda, -0.5, want, ldb,
                                   which="ref") == 0
            S = rb.DenseSkOp(rb.DenseDist(Dr, Dc, fam, ma), 3, dtype=TDT[dt])
            dA, dB = torch.from_numpy(A).to(cuda), torch.from_numpy(B0.copy()).to(cuda)
            L, O = rb.Layout(layout), rb.Op
            if side == "L":
                rb.sketch_general(L, O(opS), O(opA), p0, p1, p2, 1.5, S, off[0], off[1], dA, lda, -0.5, dB, ldb)
            else:
                rb.sketch_general(L, O(opA), O(opS), p0, p1, p2, 1.5, dA, lda, S, off[0], off[1], -0.5, dB, ldb)
            got = dB.cpu().numpy()
            # both are FP32 / FP64 GEMM-quality results of the same product (the reference's own test bound,
            # linop_common.hh:257-263, is K * 2 eps * |S||A|, looser than this)
            tol = 3e-4 if dt == "f32" else 1e-11
            assert np.allclose(got, want, rtol=tol, atol=tol * float(np.abs(want).max())), (side, layout, opS, opA, fam, K)


@needs_ref
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_sparse_operator_sketches_against_the_compiled_reference(rb, cuda, dt):
    """test_lskges.cc / test_rskges.cc shapes (SASO and LASO, offsets, transposes) vs the reference's fill_sparse + left/right_spmm"""
    rng = np.random.default_rng(6)
    npdt = np.float32 if dt == "f32" else np.float64
    cases = [("L", "C", "N", "N", 16, 20, 300, 4, "S", (0, 0)), ("L", "R", "N", "N", 64, 48, 4096, 8, "S", (0, 24)),
             ("L", "C", "N", "N", 12, 9, 200, 3, "L", (1, 2)), ("R", "C", "N", "N", 50, 10, 120, 3, "S", (0, 0)),
             ("L", "C", "T", "N", 40, 6, 10, 2, "S", (0, 0)), ("L", "C", "N", "N", 2048, 32, 8192, 8, "S", (0, 0))]
    for side, layout, opS, opA, p0, p1, p2, k, ma, off in cases:
        if side == "L":
            d, n, m = p0, p1, p2
            rS, cS = (d, m) if opS == "N" else (m, d)
            rA, cA, rB, cB = m, n, d, n
        else:
            m, d, n = p0, p1, p2
            rS, cS = (n, d) if opS == "N" else (d, n)
            rA, cA, rB, cB = m, n, m, d
        Sr, Sc = rS + off[0], cS + off[1]
        lda, ldb = (rA if layout == "C" else cA), (rB if layout == "C" else cB)
        A = rng.standard_normal(rA * cA).astype(npdt)
        B0 = rng.standard_normal(rB * cB).astype(npdt)
        want = B0.copy()
        assert ol.sketch_sparseop(dt, side, layout, opS, opA, p0, p1, p2, 1.5, Sr, Sc, k, ma, [0, 0, 0, 0], [7, 0], off[0], off[1], A, lda, -0.5, want, ldb,
                                  which="ref") == 0
        S = rb.SparseSkOp(rb.SparseDist(Sr, Sc, k, ma), 7, dtype=TDT[dt])
        dA, dB = torch.from_numpy(A).to(cuda), torch.from_numpy(B0.copy()).to(cuda)
        L, O = rb.Layout(layout), rb.Op
        if side == "L":
            rb.sketch_general(L, O(opS), O(opA), p0, p1, p2, 1.5, S, off[0], off[1], dA, lda, -0.5, dB, ldb)
        else:
            rb.sketch_general(L, O(opA), O(opS), p0, p1, p2, 1.5, dA, lda, S, off[0], off[1], -0.5, dB, ldb)
        got = dB.cpu().numpy()
        tol = 2e-5 if dt == "f32" else 1e-13
        assert np.allclose(got, want, rtol=tol, atol=tol * max(1.0, float(np.abs(want).max()))), (side, layout, opS, k, ma)


@needs_ref
def test_sketch_sparse_against_the_compiled_reference(rb, cuda):
    """lsksp3 (sksp.hh:147-192) on CSR / CSC / COO data vs the reference (test_sketch_sparse.cc is empty upstream)"""
    rng = np.random.default_rng(8)
    d, m, n, per = 48, 700, 500, 5
    colidx = np.sort(rng.integers(0, n, size=(m, per)), axis=1).astype(np.int64)
    for r in range(m):                       # make the column indices of a row distinct
        colidx[r] = np.sort(rng.choice(n, per, replace=False))
    colidx = colidx.reshape(-1)
    rowptr = np.arange(0, m * per + 1, per, dtype=np.int64)
    vals = rng.standard_normal(m * per)
    want = np.zeros(d * n)
    assert ol.sketch_sparsedata("f64", "L", "C", "N", "N", d, n, m, 1.0, d, m, "G", "L", [0, 0, 0, 0], [9, 0], 0, 0, "R", m, n, vals, rowptr, colidx,
                                0, 0, 0.0, want, d, which="ref") == 0
    A = rb.CSRMatrix(m, n, torch.from_numpy(vals).to(cuda), torch.from_numpy(rowptr).to(cuda), torch.from_numpy(colidx).to(cuda))
    S = rb.DenseSkOp(rb.DenseDist(d, m), 9, dtype=torch.float64)
    B = torch.zeros(d * n, dtype=torch.float64, device=cuda)
    rb.sketch_sparse(rb.Layout.ColMajor, rb.Op.NoTrans, rb.Op.NoTrans, d, n, m, 1.0, S, 0, 0, A, 0, 0, 0.0, B, d)
    assert np.allclose(B.cpu().numpy(), want, rtol=1e-12, atol=1e-12 * np.abs(want).max())


@needs_ref
def test_c3_tiles_against_the_compiled_reference(rb, cuda):
    """C3 (fill_dense of 2^32 floats): 64 random tiles of 2^20 entries, and the first / last tiles, bit for bit against the
    compiled reference's fill_dense of the same windows (dense_skops.hh:486-532 with (ro_s, co_s)), Gaussian and Uniform"""
    free, _ = torch.cuda.mem_get_info(cuda)
    if free < 20 * (1 << 30):
        pytest.skip("needs 18 GB of device memory")
    n, tile_cols = 65536, 16                          # natural ColMajor: 16 columns = 2^20 contiguous entries
    rng = np.random.default_rng(31)
    tiles = sorted(set([0, n // tile_cols - 1] + [int(t) for t in rng.integers(0, n // tile_cols, 64)]))
    full = torch.empty(n * n, dtype=torch.float32, device=cuda)
    for fam in ("G", "U"):
        D = rb.DenseDist(n, n, fam, "L")
        assert rb.dist_to_layout(D) == rb.Layout.ColMajor
        rb.fill_dense(D, full, rb.RNGState(11))
        for t in tiles:
            rc, exp, _ = ol.fill_dense("f32", "C", n, n, fam, "L", n, tile_cols, 0, t * tile_cols, [0, 0, 0, 0], [11, 0], which="ref")
            assert rc == 0
            got = full[t * tile_cols * n:(t + 1) * tile_cols * n].cpu().numpy()
            assert np.array_equal(got.view(np.uint32), exp.view(np.uint32)), (fam, t)
    del full


def test_boxmuller_exhaustive_against_host_libm(rb, cuda):
    """Every 32-bit raw word through the device's Box-Muller (glibc sincosf / logf restated in FP64, rbx_rng.cuh) against
    this host's libm: u0 sweeps all 2^32 values while u1 = u0 * odd + c sweeps all 2^32 values in another order, so each
    sincosf argument and each logf argument occurs exactly once (RBX_SWEEP_PAIRINGS=2 adds a second pairing, against a
    product hiding a one-ulp difference). 0 mismatches required."""
    import ctypes
    stride = int(os.environ.get("RBX_SWEEP_STRIDE", "1"))
    chunk = 1 << 25
    base = torch.arange(chunk, dtype=torch.int64, device=cuda)
    s_d = torch.empty(chunk, dtype=torch.float32, device=cuda)
    c_d = torch.empty(chunk, dtype=torch.float32, device=cuda)
    stream = ctypes.c_void_p(torch.cuda.current_stream(cuda).cuda_stream)
    bad = 0
    pairings = ((2654435761, 12345), (40503 * 65536 + 1, 0x9E3779B9))[: int(os.environ.get("RBX_SWEEP_PAIRINGS", "1"))]
    for (mul, add) in pairings:
        for c0 in range(0, 1 << 32, chunk * stride):
            u0 = ((base + c0) & 0xFFFFFFFF)
            u1 = ((u0 * mul + add) & 0xFFFFFFFF)
            u0_i = (u0 - ((u0 >> 31) << 32)).to(torch.int32)          # same 32 bits, as int32 storage
            u1_i = (u1 - ((u1 >> 31) << 32)).to(torch.int32)
            assert rb.lib.rbx_boxmuller_raw_dev(chunk, ctypes.c_void_p(u0_i.data_ptr()), ctypes.c_void_p(u1_i.data_ptr()),
                                                ctypes.c_void_p(s_d.data_ptr()), ctypes.c_void_p(c_d.data_ptr()), stream) == 0
            hs, hc = ol.boxmuller_arr(u0_i.cpu().numpy().view(np.uint32), u1_i.cpu().numpy().view(np.uint32), model=False)
            gs, gc = s_d.cpu().numpy(), c_d.cpu().numpy()
            bad += int(np.count_nonzero(gs.view(np.uint32) != hs.view(np.uint32))) + int(np.count_nonzero(gc.view(np.uint32) != hc.view(np.uint32)))
            assert bad == 0, f"device Box-Muller differs from the host libm in chunk {c0:#x} (pairing {mul:#x})"


def test_oracle_pins_on_this_box():
    """tests/test_oracle_pins.py (restatement <-> compiled reference <-> goldens, libm models vs the live libm) on the GPU
    box's own libm: the driver's `-m gpu` run would otherwise never execute it here"""
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_oracle_pins.py"), "-x", "-q", "-m", "not gpu",
                        "-p", "no:cacheprovider"], capture_output=True, text=True, timeout=1200, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
