"""GPU parity tests for operator generation (K1 fill_dense, K3 fill_sparse): BIT-EXACT against the oracle and the
golden outputs of the reference. Mirrors test/test_datastructures/test_denseskop.cc and test_sparseskop.cc."""
import json
import os

import numpy as np
import pytest
import torch

import oracle_lib as ol

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KAT = json.load(open(os.path.join(G, "philox_kat.json")))
META = json.load(open(os.path.join(G, "ref_fills.json")))
NPZ = np.load(os.path.join(G, "ref_fills.npz"))
TDT = {"f32": torch.float32, "f64": torch.float64}
IDT = {"i32": torch.int32, "i64": torch.int64}


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


def test_device_philox_kat(rb, cuda):
    import ctypes
    for v in KAT["philox4x32_10"]:
        ctr = [int(x, 16) for x in v["ctr"]]
        key = [int(x, 16) for x in v["key"]]
        out = torch.zeros(4, dtype=torch.int32, device=cuda)
        rc = rb.lib.rbx_philox4x32_dev(ol.u32(ctr), ol.u32(key), 1, ctypes.c_void_p(out.data_ptr()), None)
        assert rc == 0
        assert [int(x) & 0xFFFFFFFF for x in out.cpu().tolist()] == [int(x, 16) for x in v["out"]]
    # stream with carries: ctr + i for a counter that overflows words 0 and 1
    ctr, key, n = [0xFFFFFFF0, 0xFFFFFFFF, 7, 0], [1, 2], 64
    out = torch.zeros(4 * n, dtype=torch.int32, device=cuda)
    assert rb.lib.rbx_philox4x32_dev(ol.u32(ctr), ol.u32(key), n, ctypes.c_void_p(out.data_ptr()), None) == 0
    got = out.cpu().numpy().view(np.uint32).reshape(n, 4)
    for i in range(n):
        assert got[i].tolist() == ol.philox(ol.ctr_add(ctr, i), key)


def test_boxmuller_transform_bits_vs_host_libm(rb, cuda):
    """the device transform equals r123::boxmuller on the host libm for 4M random word pairs + edge words"""
    import ctypes
    rng = np.random.default_rng(7)
    n = 1 << 22
    u0 = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    u1 = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    edge = np.array([0, 1, 2, 3, 0x7FFFFFFF, 0x80000000, 0x80000001, 0xFFFFFFFF, 0xFFFFFFFE, 0x3FFFFFFF, 0x40000000, 0xC0000000,
                     0x00000800, 0x000007FF, 0xFFFFF800, 0x20000000, 0x5FFFFFFF, 0xA0000000], np.uint32)
    e0, e1 = np.meshgrid(edge, edge)
    u0[: e0.size], u1[: e1.size] = e0.ravel(), e1.ravel()
    # a dense run of consecutive words around the quadrant boundaries of sincos and the table edges of log
    u0[e0.size: e0.size + 4096] = np.arange(0x3FFFF800, 0x3FFFF800 + 4096, dtype=np.uint64).astype(np.uint32)
    d0 = torch.from_numpy(u0.view(np.int32)).to(cuda)
    d1 = torch.from_numpy(u1.view(np.int32)).to(cuda)
    s = torch.empty(n, dtype=torch.float32, device=cuda)
    c = torch.empty(n, dtype=torch.float32, device=cuda)
    vp = ctypes.c_void_p
    assert rb.lib.rbx_boxmuller_raw_dev(n, vp(d0.data_ptr()), vp(d1.data_ptr()), vp(s.data_ptr()), vp(c.data_ptr()), None) == 0
    hs, hc = ol.boxmuller_arr(u0, u1, model=False)
    assert np.array_equal(bits(s.cpu().numpy()), bits(hs))
    assert np.array_equal(bits(c.cpu().numpy()), bits(hc))


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "fill_dense"], ids=lambda m: m["id"])
def test_fill_dense_vs_reference_golden(rb, cuda, m):
    D = rb.DenseDist(m["D_rows"], m["D_cols"], m["family"], m["major_axis"])
    buf = torch.full((m["n_rows"] * m["n_cols"],), float("nan"), dtype=TDT[m["dt"]], device=cuda)
    nxt = rb.fill_dense(rb.Layout(m["layout"]), D, m["n_rows"], m["n_cols"], m["ro_s"], m["co_s"], buf, rb.RNGState(m["key"], m["ctr"]))
    if m["family"] == "G":
        # Gaussian goldens carry the authoring host's libm bits; compare with this host's libm through the oracle too
        rc, exp, _ = ol.fill_dense(m["dt"], m["layout"], m["D_rows"], m["D_cols"], m["family"], m["major_axis"], m["n_rows"], m["n_cols"],
                                   m["ro_s"], m["co_s"], m["ctr"], m["key"])
        assert np.array_equal(bits(buf.cpu().numpy()), bits(exp))
    assert np.array_equal(bits(buf.cpu().numpy()), NPZ[m["id"]])
    assert nxt.counter == m["next"]


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "submat"], ids=lambda m: m["id"])
def test_fill_dense_submat_vs_reference_golden(rb, cuda, m):
    ld = m["lda"] if m["lda"] > 0 else m["n_scols"]
    buf = torch.full((m["n_srows"] * ld,), float("nan"), dtype=torch.float32, device=cuda)
    nxt = rb.fill_dense_submat(m["family"], m["n_cols"], buf, m["n_srows"], m["n_scols"], m["ptr"], rb.RNGState(m["key"], m["ctr"]), m["lda"])
    got = bits(buf.cpu().numpy()).reshape(m["n_srows"], ld)
    exp = NPZ[m["id"]].reshape(m["n_srows"], ld)
    assert np.array_equal(got[:, : m["n_scols"]], exp[:, : m["n_scols"]])
    # padding between rows (lda > n_scols) must be left untouched
    assert np.all(np.isnan(buf.cpu().numpy().reshape(m["n_srows"], ld)[:, m["n_scols"]:]))
    assert nxt.counter == m["next"]


@pytest.mark.parametrize("fam", ["G", "U"])
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_fill_dense_random_windows_vs_oracle(rb, cuda, fam, dt):
    rng = np.random.default_rng(11)
    for _ in range(40):
        R, C = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        nr, nc = int(rng.integers(1, R + 1)), int(rng.integers(1, C + 1))
        ro, co = int(rng.integers(0, R - nr + 1)), int(rng.integers(0, C - nc + 1))
        ma, lay = "LS"[rng.integers(2)], "CR"[rng.integers(2)]
        ctr = [int(rng.integers(0, 2**32)), 0xFFFFFFFF * int(rng.integers(0, 2)), int(rng.integers(0, 3)), 0]
        key = [int(rng.integers(0, 2**32)), int(rng.integers(0, 2**32))]
        buf = torch.empty(nr * nc, dtype=TDT[dt], device=cuda)
        nxt = rb.fill_dense(rb.Layout(lay), rb.DenseDist(R, C, fam, ma), nr, nc, ro, co, buf, rb.RNGState(key, ctr))
        rc, exp, enx = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key)
        assert rc == 0 and np.array_equal(bits(buf.cpu().numpy()), bits(exp)), (R, C, nr, nc, ro, co, ma, lay)
        assert nxt.counter == enx


def test_submatrix_consistency_and_state_chaining(rb, cuda):
    """test_denseskop.cc:162-296 (windows of a full fill) and :405-440 (next_state chaining == concatenation)."""
    for fam in "GU":
        D = rb.DenseDist(37, 203, fam, "L")
        full = torch.empty(37 * 203, dtype=torch.float32, device=cuda)
        rb.fill_dense(D, full, rb.RNGState(3))
        F = full.view(37, 203)
        for (nr, nc, ro, co) in [(5, 7, 3, 1), (37, 1, 0, 202), (1, 203, 36, 0), (10, 100, 20, 99), (4, 4, 33, 199)]:
            w = torch.empty(nr * nc, dtype=torch.float32, device=cuda)
            rb.fill_dense(rb.Layout.RowMajor, D, nr, nc, ro, co, w, rb.RNGState(3))
            assert torch.equal(w.view(nr, nc), F[ro:ro + nr, co:co + nc])
            wt = torch.empty(nr * nc, dtype=torch.float32, device=cuda)
            rb.fill_dense(rb.Layout.ColMajor, D, nr, nc, ro, co, wt, rb.RNGState(3))
            assert torch.equal(wt.view(nc, nr).T, F[ro:ro + nr, co:co + nc])
        # chaining: S2 seeded with S1.next_state continues the stream as if one taller short-axis-major operator
        D1 = rb.DenseDist(64, 16, fam, "S")           # tall, Short-axis major => RowMajor, rows are counter rows
        S1 = rb.DenseSkOp(D1, 9)
        S2 = rb.DenseSkOp(D1, S1.next_state)
        D12 = rb.DenseDist(128, 16, fam, "S")
        b1 = torch.empty(64 * 16, device=cuda)
        b2 = torch.empty(64 * 16, device=cuda)
        b12 = torch.empty(128 * 16, device=cuda)
        n1 = rb.fill_dense(D1, b1, S1.seed_state)
        rb.fill_dense(D1, b2, S2.seed_state)
        rb.fill_dense(D12, b12, rb.RNGState(9))
        assert n1 == S1.next_state
        assert torch.equal(torch.cat([b1, b2]), b12)


def test_fill_dense_large_counter_offsets(rb, cuda):
    """config-5-sized parent (8192 x 2^23, W = 2^21): windows whose counters exceed 2^32 (H9)"""
    D = rb.DenseDist(8192, 1 << 23, "G", "L")
    for (ro, co) in [(8191, (1 << 23) - 64), (2049, 4194301), (4096, 0)]:
        w = torch.empty(2 * 64, dtype=torch.float64, device=cuda)
        nr = min(2, 8192 - ro)
        nc = min(64, (1 << 23) - co)
        rb.fill_dense(rb.Layout.RowMajor, D, nr, nc, ro, co, w[: nr * nc], rb.RNGState(0))
        rc, exp, _ = ol.fill_dense("f64", "R", 8192, 1 << 23, "G", "L", nr, nc, ro, co, [0, 0, 0, 0], [0, 0])
        assert np.array_equal(bits(w[: nr * nc].cpu().numpy()), bits(exp))


@pytest.mark.parametrize("fam", ["G", "U"])
def test_fill_dense_seed_counters_that_carry_inside_the_fill(rb, cuda, fam):
    """whole-matrix fills (the linear loop) from seed states whose counter words overflow in the middle of the fill: word 0
    into word 1 (running 64-bit sum), nonzero high words (loop constants), and the low 64 bits wrapping into word 2 (the loop
    with the full 128-bit add); both output precisions; the returned next state carries too (base.hh:215-232)"""
    n_rows, n_cols = 96, 1000                       # 24 000 counters
    for ctr in ([0xFFFFFF00, 0, 0, 0], [0xFFFFF000, 0xFFFFFFFF, 5, 9], [0xFFFFC000, 0xFFFFFFFF, 0xFFFFFFFF, 3], [7, 1, 0xFFFFFFFF, 0xFFFFFFFF]):
        for dt, tdt in (("f32", torch.float32), ("f64", torch.float64)):
            buf = torch.empty(n_rows * n_cols, dtype=tdt, device=cuda)
            nxt = rb.fill_dense(rb.DenseDist(n_rows, n_cols, fam), buf, rb.RNGState(11, ctr))
            rc, exp, enxt = ol.fill_dense(dt, "R", n_rows, n_cols, fam, "L", n_rows, n_cols, 0, 0, ctr, [11, 0])
            assert rc == 0 and np.array_equal(bits(buf.cpu().numpy()), bits(exp)), (fam, dt, ctr)
            assert list(nxt.counter) == list(enxt)


def test_fill_dense_2pow28_tiles_vs_oracle(rb, cuda):
    """a 2^28-entry Gaussian fill (16384 x 16384): full-buffer determinism + 32 random 64x64 tiles bitwise vs the oracle"""
    n = 16384
    D = rb.DenseDist(n, n, "G", "L")
    a = torch.empty(n * n, dtype=torch.float32, device=cuda)
    b = torch.empty(n * n, dtype=torch.float32, device=cuda)
    rb.fill_dense(D, a, rb.RNGState(0))
    rb.fill_dense(D, b, rb.RNGState(0))
    assert torch.equal(a.view(torch.int32), b.view(torch.int32))
    lay = rb.dist_to_layout(D).value          # square => ColMajor natural
    M = a.view(n, n).T if lay == "C" else a.view(n, n)
    rng = np.random.default_rng(5)
    for _ in range(32):
        ro, co = int(rng.integers(0, n - 64)), int(rng.integers(0, n - 64))
        rc, exp, _ = ol.fill_dense("f32", "R", n, n, "G", "L", 64, 64, ro, co, [0, 0, 0, 0], [0, 0])
        got = M[ro:ro + 64, co:co + 64].contiguous().cpu().numpy()
        assert np.array_equal(bits(got).ravel(), bits(exp))
    assert bool(torch.isfinite(a).all())
    assert abs(float(a.mean())) < 1e-3 and abs(float(a.std()) - 1.0) < 1e-3   # test_denseskop.cc:137-159


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "fill_sparse"], ids=lambda m: m["id"])
def test_fill_sparse_vs_reference_golden(rb, cuda, m):
    S = rb.SparseSkOp(rb.SparseDist(m["n_rows"], m["n_cols"], m["vec_nnz"], m["major_axis"]), rb.RNGState(m["key"], m["ctr"]),
                      dtype=TDT[m["dt"]], index_dtype=IDT[m["it"]])
    rb.fill_sparse(S)
    assert np.array_equal(S.rows.cpu().numpy(), NPZ[m["id"] + "_rows"])
    assert np.array_equal(S.cols.cpu().numpy(), NPZ[m["id"] + "_cols"])
    assert np.array_equal(S.vals.cpu().numpy(), NPZ[m["id"] + "_vals"])
    assert S.next_state.counter == m["next"]


def test_fill_sparse_random_vs_oracle(rb, cuda):
    rng = np.random.default_rng(3)
    cases = [(2048, 65536, 8, "S"), (65536, 2048, 8, "S"), (300, 5000, 100, "S"), (300, 5000, 7, "L"), (5000, 300, 150, "L"),
             (40, 200, 40, "S"), (1, 50, 1, "S"), (50, 1, 1, "L")]
    for (r, c, k, ma) in cases:
        key = [int(rng.integers(0, 2**32)), 0]
        ctr = [0xFFFFFF00, 0xFFFFFFFF, 0, 0]
        for dt, it in (("f64", "i64"), ("f32", "i32")):
            S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), rb.RNGState(key, ctr), dtype=TDT[dt], index_dtype=IDT[it])
            rb.fill_sparse(S)
            rc, rows, cols, vals, nxt = ol.fill_sparse(dt, it, r, c, k, ma, ctr, key)
            assert rc == 0
            assert np.array_equal(S.rows.cpu().numpy(), rows) and np.array_equal(S.cols.cpu().numpy(), cols)
            assert np.array_equal(S.vals.cpu().numpy(), vals)
            assert S.next_state.counter == nxt


def test_fill_sparse_long_vectors_hash_kernel(rb, cuda, monkeypatch):
    """16 < vec_nnz <= 8192: one CTA per vector with the lazily kept work array in shared memory (direct array below vec_nnz, hash
    table above), 128 shuffle steps at a time resolved in rounds of steps that do not interact. Cases where almost no steps
    interact (vec_nnz << the sampled axis), where almost all do (a full permutation of the axis), several groups of steps, the
    largest table, and one step beyond it (the global work-array kernel); always the oracle's indices and signs, and the same
    as with the kernel off."""
    cases = [(300, 5000, 300, "S"), (3000, 3000, 3000, "S"), (64, 50000, 4096, "L"), (200, 9000, 8192, "L"), (9000, 150, 8193, "L"),
             (2048, 40000, 33, "S"), (5000, 700, 513, "S"), (100, 1 << 20, 1000, "L"), (64, 70000, 17, "S"), (20, 3000, 20, "S"), (129, 129, 129, "S")]
    for (r, c, k, ma) in cases:
        key, ctr = [k * 7 + r, 1], [0xFFFFFFF0, 0xFFFFFFFF, 3, 0]
        S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), rb.RNGState(key, ctr), dtype=torch.float32, index_dtype=torch.int32)
        rb.fill_sparse(S)
        rc, rows, cols, vals, nxt = ol.fill_sparse("f32", "i32", r, c, k, ma, ctr, key)
        assert rc == 0
        assert np.array_equal(S.rows.cpu().numpy(), rows) and np.array_equal(S.cols.cpu().numpy(), cols), (r, c, k, ma)
        assert np.array_equal(S.vals.cpu().numpy(), vals) and S.next_state.counter == nxt
    monkeypatch.setenv("RBX_FY_NO_HASH", "1")
    for (r, c, k, ma) in cases[:3]:
        S = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), 5); rb.fill_sparse(S)
        monkeypatch.delenv("RBX_FY_NO_HASH")
        S2 = rb.SparseSkOp(rb.SparseDist(r, c, k, ma), 5); rb.fill_sparse(S2)
        monkeypatch.setenv("RBX_FY_NO_HASH", "1")
        assert torch.equal(S.rows, S2.rows) and torch.equal(S.cols, S2.cols) and torch.equal(S.vals, S2.vals)


def test_repeated_fisher_yates_workarray_path(rb, cuda):
    """repeated_fisher_yates as its own entry point (vec_nnz = 200: the per-CTA hash kernel); same stream as the oracle"""
    k, n, r = 200, 1000, 37
    idx = torch.empty(k * r, dtype=torch.int64, device=cuda)
    rb.repeated_fisher_yates(rb.RNGState(7), k, n, r, idx)
    rc, rows, cols, vals, _ = ol.fill_sparse("f64", "i64", n, r, k, "L", [0, 0, 0, 0], [7, 0])   # tall LASO: F-Y over the long axis per column
    assert rc == 0 and np.array_equal(idx.cpu().numpy(), rows)
    for v in idx.view(r, k).cpu().numpy():
        assert len(set(v.tolist())) == k


# ---- index sampling utilities (util.hh:257-334): bit-exact against the oracle ------------------------------------
@pytest.mark.parametrize("n,k", [(10, 0), (10, 1), (7, 4), (1000, 5), (1 << 20, 100003), (3, 64)])
def test_sample_indices_iid_uniform(rb, cuda, n, k):
    ctr, key = [0xFFFFFFFE, 0xFFFFFFFF, 0, 0], [42, 0]
    out = torch.full((max(k, 1),), -7, dtype=torch.int64, device=cuda)
    nxt = rb.sample_indices_iid_uniform(n, k, out, rb.RNGState(key, ctr))
    rc, want, nx = ol.sample_indices_iid_uniform(n, k, ctr, key)
    assert rc == 0 and np.array_equal(out.cpu().numpy()[:k], want) and list(nxt.counter) == nx
    if k == 0:
        assert int(out[0]) == -7


@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_weights_to_cdf_and_sample_indices_iid(rb, cuda, dt):
    tdt = torch.float32 if dt == "f32" else torch.float64
    rng = np.random.default_rng(3)
    for n, k in [(1, 5), (17, 40), (1000, 1003), (100000, 20001)]:
        w = rng.random(n)
        w[rng.random(n) < 0.2] = 0.0
        dw = torch.from_numpy(w).to(tdt).to(cuda)
        rb.weights_to_cdf(n, dw)
        rc, cdf = ol.weights_to_cdf(dt, w)
        assert rc == 0 and np.array_equal(dw.cpu().numpy().view(np.uint8), cdf.view(np.uint8))      # sequential sum: same bits
        out = torch.zeros(k, dtype=torch.int64, device=cuda)
        nxt = rb.sample_indices_iid(n, dw, k, out, rb.RNGState([7, 1], [5, 0, 0, 0]))
        rc, want, nx = ol.sample_indices_iid(dt, cdf, k, [5, 0, 0, 0], [7, 1])
        assert rc == 0 and np.array_equal(out.cpu().numpy(), want) and list(nxt.counter) == nx
    for bad in ([0.5, -0.1, 0.5], [0.0, 0.0, 0.0]):
        with pytest.raises(rb.Error):
            rb.weights_to_cdf(3, torch.tensor(bad, dtype=tdt, device=cuda))
