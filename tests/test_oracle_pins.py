"""CPU tests that PIN the oracle (oracle/randblas_oracle.c) before anything trusts it:
   (1) the reference's own known-answer vectors and counter-carry cases,
   (2) golden outputs of the reference itself (tests/golden/ref_fills.*, made by make_golden.py),
   (3) when oracle/_ref is built, live comparison with the compiled reference on fresh random cases,
   (4) the explicit-FMA models of glibc sincosf/logf against the live host libm.
"""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KAT = json.load(open(os.path.join(G, "philox_kat.json")))
META = json.load(open(os.path.join(G, "ref_fills.json")))
NPZ = np.load(os.path.join(G, "ref_fills.npz"))


def bits(a):
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64)


def test_philox_kat():
    assert len(KAT["philox4x32_10"]) == 3
    for v in KAT["philox4x32_10"]:
        ctr = [int(x, 16) for x in v["ctr"]]
        key = [int(x, 16) for x in v["key"]]
        assert ol.philox(ctr, key) == [int(x, 16) for x in v["out"]]
        if ol.REF is not None:
            assert ol.philox(ctr, key, "ref") == [int(x, 16) for x in v["out"]]


def test_counter_carry():
    for case in KAT["incr"]:
        for which in (["orc", "ref"] if ol.REF is not None else ["orc"]):
            c = list(case["start"])
            for n in case["incr"]:
                c = ol.ctr_add(c, n, which)
            assert c == case["expect"], (case, which)


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "fill_dense"][::3], ids=lambda m: m["id"])
def test_fill_dense_vs_golden(m):
    rc, out, nxt = ol.fill_dense(m["dt"], m["layout"], m["D_rows"], m["D_cols"], m["family"], m["major_axis"], m["n_rows"], m["n_cols"],
                                 m["ro_s"], m["co_s"], m["ctr"], m["key"])
    assert rc == 0
    assert np.array_equal(bits(out), NPZ[m["id"]])
    assert nxt == m["next"]


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "submat"], ids=lambda m: m["id"])
def test_submat_vs_golden(m):
    rc, out, nxt = ol.fill_dense_submat("f32", m["family"], m["n_cols"], m["n_srows"], m["n_scols"], m["ptr"], m["ctr"], m["key"], m["lda"])
    assert rc == 0
    ld = m["lda"] if m["lda"] > 0 else m["n_scols"]
    got = bits(out).reshape(m["n_srows"], ld)[:, : m["n_scols"]]
    exp = NPZ[m["id"]].reshape(m["n_srows"], ld)[:, : m["n_scols"]]
    assert np.array_equal(got, exp)
    assert nxt == m["next"]


@pytest.mark.parametrize("m", [m for m in META if m["kind"] == "fill_sparse"][::2], ids=lambda m: m["id"])
def test_fill_sparse_vs_golden(m):
    rc, rows, cols, vals, nxt = ol.fill_sparse(m["dt"], m["it"], m["n_rows"], m["n_cols"], m["vec_nnz"], m["major_axis"], m["ctr"], m["key"])
    assert rc == 0
    assert np.array_equal(rows, NPZ[m["id"] + "_rows"])
    assert np.array_equal(cols, NPZ[m["id"] + "_cols"])
    assert np.array_equal(vals, NPZ[m["id"] + "_vals"])
    assert nxt == m["next"]
    # the reference's own property (test_sparseskop.cc:48-104): no duplicate index inside a major-axis vector
    k = m["vec_nnz"]
    lng_is_cols = m["n_cols"] >= m["n_rows"]
    short_idx = rows if (m["n_rows"] == min(m["n_rows"], m["n_cols"])) else cols
    long_idx = cols if short_idx is rows else rows
    major = short_idx if m["major_axis"] == "S" else long_idx
    for v in major.reshape(-1, k):
        assert len(set(v.tolist())) == k
    del lng_is_cols


@pytest.mark.skipif(ol.REF is None, reason="compiled reference (oracle/_ref) not built on this machine")
def test_oracle_vs_compiled_reference_random():
    rng = np.random.default_rng(0)
    for _ in range(60):
        R, C = int(rng.integers(1, 40)), int(rng.integers(1, 40))
        nr, nc = int(rng.integers(1, R + 1)), int(rng.integers(1, C + 1))
        ro, co = int(rng.integers(0, R - nr + 1)), int(rng.integers(0, C - nc + 1))
        fam, ma, lay, dt = "GU"[rng.integers(2)], "LS"[rng.integers(2)], "CR"[rng.integers(2)], ("f32", "f64")[rng.integers(2)]
        ctr = [int(rng.integers(0, 2**32)), int(rng.integers(0, 2)) * 0xFFFFFFFF, 0, 0]
        key = [int(rng.integers(0, 2**32)), int(rng.integers(0, 2**32))]
        a = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key, "orc")
        b = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key, "ref")
        assert a[0] == b[0] == 0 and np.array_equal(bits(a[1]), bits(b[1])) and a[2] == b[2]
    for _ in range(40):
        r, c = int(rng.integers(1, 50)), int(rng.integers(1, 50))
        ma = "SL"[rng.integers(2)]
        k = int(rng.integers(1, (min(r, c) if ma == "S" else max(r, c)) + 1))
        key = [int(rng.integers(0, 2**32)), 0]
        a = ol.fill_sparse("f64", "i64", r, c, k, ma, [9, 0, 0, 0], key, "orc")
        b = ol.fill_sparse("f64", "i64", r, c, k, ma, [9, 0, 0, 0], key, "ref")
        for x, y in zip(a[1:4], b[1:4]):
            assert np.array_equal(x, y)
        assert a[4] == b[4]


@pytest.mark.skipif(ol.REF is None, reason="compiled reference (oracle/_ref) not built on this machine")
def test_oracle_vs_compiled_reference_random_larger():
    """the same comparison at the sizes the randomized GPU tests draw from (tests/test_gpu_fuzz.py checks the device against the
    oracle): long Fisher-Yates vectors (vec_nnz up to 700, full permutations included), both major axes, wide and tall; dense
    windows of operators with thousands of rows / columns and counters about to carry"""
    rng = np.random.default_rng(17)
    for _ in range(24):
        r, c = int(rng.integers(1, 3000)), int(rng.integers(1, 3000))
        ma = "SL"[rng.integers(2)]
        k = int(min(r, c, [1, 8, 17, 33, 129, 700][rng.integers(6)]))
        if rng.random() < 0.2:
            k = min(r, c)                       # a full permutation of the short axis
        while k * (max(r, c) if ma == "S" else min(r, c)) > 1500000:
            k = max(1, k // 2)
        key, ctr = [int(rng.integers(0, 2**32)), int(rng.integers(0, 2**32))], [0xFFFFFF00, 0xFFFFFFFF, int(rng.integers(0, 2**32)), 0]
        it = ("i32", "i64")[rng.integers(2)]
        a = ol.fill_sparse("f32", it, r, c, k, ma, ctr, key, "orc")
        b = ol.fill_sparse("f32", it, r, c, k, ma, ctr, key, "ref")
        assert a[0] == b[0] == 0
        for x, y in zip(a[1:4], b[1:4]):
            assert np.array_equal(x, y), (r, c, k, ma)
        assert a[4] == b[4]
    for _ in range(24):
        R, C = int(rng.integers(1, 3000)), int(rng.integers(1, 3000))
        nr, nc = int(rng.integers(1, min(R, 300) + 1)), int(rng.integers(1, min(C, 300) + 1))
        ro, co = int(rng.integers(0, R - nr + 1)), int(rng.integers(0, C - nc + 1))
        fam, ma, lay, dt = "GU"[rng.integers(2)], "LS"[rng.integers(2)], "CR"[rng.integers(2)], ("f32", "f64")[rng.integers(2)]
        ctr = [0xFFFFFFFF - int(rng.integers(0, 100000)), 0xFFFFFFFF, int(rng.integers(0, 2**32)), 0]
        key = [int(rng.integers(0, 2**32)), int(rng.integers(0, 2**32))]
        a = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key, "orc")
        b = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key, "ref")
        assert a[0] == b[0] == 0 and np.array_equal(bits(a[1]), bits(b[1])) and a[2] == b[2], (R, C, nr, nc, ro, co, fam, ma, lay, dt)


def test_uniform_transforms_documented_properties():
    """Random123's u01 / uneg11 (uniform.hpp) are restated in oracle/shims (no checkout offline). Their documented contract
    pins the restatement at the edges: u01<float> maps into (0, 1] -- never 0, exactly 1 at the top word -- and is
    monotone; uneg11<float> maps into [-1, 1], never 0, antisymmetric about the half-way word up to the 2^-32 offset."""
    u01, un = ol.ORC.orc_u01, ol.ORC.orc_uneg11
    assert u01(0) == 2.0 ** -33 and u01(0xFFFFFFFF) == 1.0 and u01(0x80000000) == 0.5
    assert un(0) == 2.0 ** -32 and un(0x7FFFFFFF) == 1.0 and un(0x80000000) == -1.0 and un(0xFFFFFFFF) == np.float32(-2.0 ** -31 + 2.0 ** -32)
    words = np.array([0, 1, 2, 255, 256, 0x7FFFFF, 0x800000, 0x1000000, 0x7FFFFFFF, 0x80000000, 0x80000001, 0xFFFFFF00, 0xFFFFFFFF], np.uint64)
    vals = [u01(int(w)) for w in words]
    assert all(0.0 < v <= 1.0 for v in vals) and all(a <= b for a, b in zip(vals, vals[1:]))
    assert all(-1.0 <= un(int(w)) <= 1.0 and un(int(w)) != 0.0 for w in words)
    # the compiled reference (through the shims) produces the same floats lane by lane
    if ol.REF is not None:
        for ctr in ([0, 0, 0, 0], [5, 0, 0, 7]):
            raw = ol.philox(ctr, [3, 0], "ref")
            got = ol.generate4("U", ctr, [3, 0], "ref")
            assert [float(g) for g in got] == [un(w) for w in raw]


def test_libm_models_match_host_libm_strided():
    """oracle's explicit-FMA models == this host's sincosf/logf on a 2^22-point strided sweep of the 2^32 domain
    (the full sweep, run once by hand, is recorded in DESIGN.md: 0 mismatches on glibc 2.39 / FMA)."""
    import ctypes
    fb = ctypes.c_uint32(0)
    assert ol.ORC.orc_sweep_sincos(0, 1 << 32, 1021, ctypes.byref(fb)) == 0, hex(fb.value)
    assert ol.ORC.orc_sweep_log(0, 1 << 32, 1021, ctypes.byref(fb)) == 0, hex(fb.value)
    # edge words: u01 -> 1.0f (log = 0), uneg11 -> +-1 (sincos(+-pi)), smallest magnitudes
    edge = np.array([0, 1, 2, 0x7FFFFFFF, 0x80000000, 0x80000001, 0xFFFFFFFF, 0xFFFFFFFE, 0x3FFFFFFF, 0x40000000, 0xC0000000], np.uint32)
    u0, u1 = np.meshgrid(edge, edge)
    s0, c0 = ol.boxmuller_arr(u0.ravel(), u1.ravel(), model=False)
    s1, c1 = ol.boxmuller_arr(u0.ravel(), u1.ravel(), model=True)
    assert np.array_equal(s0.view(np.uint32), s1.view(np.uint32)) and np.array_equal(c0.view(np.uint32), c1.view(np.uint32))


@pytest.mark.skipif(ol.REF is None, reason="compiled reference (oracle/_ref) not built on this machine")
def test_oracle_sketch_definitions_vs_reference():
    """the oracle's definitional sketches agree with the reference's lskge3/rskge3/lskges within the reference's own bound"""
    rng = np.random.default_rng(1)
    for dt, eps in (("f32", 2.0**-23), ("f64", 2.0**-52)):
        npdt = np.float32 if dt == "f32" else np.float64
        for side in "LR":
            for lay in "CR":
                for opS in "NT":
                    d, n, m = 5, 6, 17
                    A = rng.standard_normal(m * n).astype(npdt)
                    p = (d, n, m) if side == "L" else (n, d, m)   # right: B(n' x d) = A(n' x m) S(m x d); reuse sizes
                    rowsS, colsS = ((d, m) if opS == "N" else (m, d)) if side == "L" else ((m, d) if opS == "N" else (d, m))
                    Dr, Dc = rowsS + 2, colsS + 3
                    lda = (m if lay == "C" else n) if side == "L" else (n if lay == "C" else m)
                    ldb = (d if lay == "C" else n) if side == "L" else (n if lay == "C" else d)
                    B0 = rng.standard_normal(d * n).astype(npdt)
                    Ba, Bb = B0.copy(), B0.copy()
                    args = (dt, side, lay, opS, "N", *p, 1.5, Dr, Dc, "G", "L", [0, 0, 0, 0], [3, 0], 2, 3, A, lda, -0.5)
                    assert ol.sketch_dense(*args, Ba, ldb, "orc") == 0
                    assert ol.sketch_dense(*args, Bb, ldb, "ref") == 0
                    assert np.allclose(Ba, Bb, rtol=0, atol=1.5 * m * 2 * eps * 40 + 0.5 * eps * 4), (dt, side, lay, opS)


# ---- index sampling utilities: restatement vs the compiled reference (util.hh:257-334) --------------------------
@pytest.mark.skipif(ol.REF is None, reason="compiled reference not built")
@pytest.mark.parametrize("n,k", [(10, 0), (10, 1), (7, 4), (1000, 5), (1 << 20, 1001), (3, 64)])
def test_sample_indices_iid_uniform_matches_reference(n, k):
    ctr, key = [0xFFFFFFFE, 0xFFFFFFFF, 0, 0], [42, 0]          # counter carries into the third word
    rc_o, s_o, nx_o = ol.sample_indices_iid_uniform(n, k, ctr, key, "orc")
    rc_r, s_r, nx_r = ol.sample_indices_iid_uniform(n, k, ctr, key, "ref")
    assert rc_o == rc_r == 0 and np.array_equal(s_o, s_r) and nx_o == nx_r
    assert (s_o >= 0).all() and (s_o <= n).all()


@pytest.mark.skipif(ol.REF is None, reason="compiled reference not built")
@pytest.mark.parametrize("dt", ["f32", "f64"])
def test_weights_to_cdf_and_sample_indices_iid_match_reference(dt):
    rng = np.random.default_rng(3)
    for n, k in [(1, 5), (17, 40), (1000, 1003)]:
        w = rng.random(n)
        w[rng.random(n) < 0.2] = 0.0                                # flat stretches in the cdf
        rc_o, cdf_o = ol.weights_to_cdf(dt, w, which="orc")
        rc_r, cdf_r = ol.weights_to_cdf(dt, w, which="ref")
        assert rc_o == rc_r == 0 and np.array_equal(cdf_o.view(np.uint8), cdf_r.view(np.uint8))
        rc_o, s_o, nx_o = ol.sample_indices_iid(dt, cdf_o, k, [5, 0, 0, 0], [7, 1], "orc")
        rc_r, s_r, nx_r = ol.sample_indices_iid(dt, cdf_o, k, [5, 0, 0, 0], [7, 1], "ref")
        assert rc_o == rc_r == 0 and np.array_equal(s_o, s_r) and nx_o == nx_r
    # the reference throws on a weight below the threshold and on an all-zero vector
    for bad in ([0.5, -0.1, 0.5], [0.0, 0.0, 0.0]):
        assert ol.weights_to_cdf(dt, bad, which="orc")[0] != 0 and ol.weights_to_cdf(dt, bad, which="ref")[0] != 0
