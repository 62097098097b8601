"""Generates tests/golden/*.json|npz from the REFERENCE itself (run in the authoring container only).

Sources:
  * philox_kat.json  : copied numbers from /root/reference/test/test_basic_rng/r123_kat_vectors.txt:19-21
                       (Philox4x32-10 known answers) and the counter-carry cases of test_r123.cc:679-766.
  * ref_fills.npz    : outputs of the unmodified reference headers (oracle/_ref/librandblas_ref.so, built by
                       oracle/Makefile from /root/reference) for fill_dense / fill_dense_submat_impl / fill_sparse
                       on the shapes the reference's own tests use (test_denseskop.cc, test_sparseskop.cc).
Gaussian entries depend on the host libm (random_gen.hh:61-65 uses sincosf/logf); the container's glibc is 2.39
with FMA variants selected, which is what the CUDA kernel models. The raw bit patterns are stored (uint32/uint64).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

REFROOT = "/root/reference"


def kat():
    out = {"philox4x32_10": [], "incr": []}
    for line in open(os.path.join(REFROOT, "test/test_basic_rng/r123_kat_vectors.txt")):
        t = line.split()
        if len(t) == 12 and t[0] == "philox4x32" and t[1] == "10":
            out["philox4x32_10"].append({"ctr": t[2:6], "key": t[6:8], "out": t[8:12]})
    m = 0xFFFFFFFF
    # (start counter, sequence of increments, expected counter) -- test_r123.cc:679-766
    out["incr"] = [
        {"start": [0, 0, 0, 0], "incr": [m], "expect": [m, 0, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [m, 1], "expect": [0, 1, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [m, 1, 3], "expect": [3, 1, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [(1 << 32) - 1], "expect": [m, 0, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [1 << 32], "expect": [0, 1, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [1 << 63, (1 << 63) - (1 << 32)], "expect": [0, m, 0, 0]},
        {"start": [0, 0, 0, 0], "incr": [1 << 63, (1 << 63) - (1 << 32), 1 << 32], "expect": [0, 0, 1, 0]},
        {"start": [m, m, m, 0], "incr": [1], "expect": [0, 0, 0, 1]},
    ]
    json.dump(out, open(os.path.join(HERE, "philox_kat.json"), "w"), indent=1)


def bits(a):
    return a.view(np.uint32 if a.dtype == np.float32 else np.uint64).copy()


def fills():
    z = {}
    meta = []
    # fill_dense on the reference's test shapes (test_denseskop.cc) + lane/padding phases
    cases = []
    for fam in "GU":
        for dt in ("f32", "f64"):
            for (r, c) in [(7, 20), (20, 7), (8, 8), (3, 10), (10, 3), (5, 13), (1, 9), (9, 1)]:
                for ma in "LS":
                    cases.append((fam, dt, r, c, ma, r, c, 0, 0, None))
            # submatrices with offsets, incl. n_cols % 4 in {0,1,2,3} and unaligned starts
            for (R, C, nr, nc, ro, co) in [(8, 12, 5, 7, 3, 1), (12, 8, 7, 5, 1, 3), (10, 13, 4, 9, 2, 3), (13, 10, 9, 4, 3, 2),
                                           (6, 15, 6, 6, 0, 5), (9, 9, 4, 4, 5, 5)]:
                for ma in "LS":
                    for lay in "CR":
                        cases.append((fam, dt, R, C, ma, nr, nc, ro, co, lay))
    for k, (fam, dt, R, C, ma, nr, nc, ro, co, lay) in enumerate(cases):
        key = [42 + k % 3, 0]
        ctr = [k * 7 % 5, 0, 0, 0]
        if lay is None:
            lay = "R" if (R < C) == (ma == "L") else "C"
        rc, out, nxt = ol.fill_dense(dt, lay, R, C, fam, ma, nr, nc, ro, co, ctr, key, which="ref")
        assert rc == 0
        z[f"fd{k}"] = bits(out)
        meta.append({"kind": "fill_dense", "id": f"fd{k}", "dt": dt, "layout": lay, "D_rows": R, "D_cols": C, "family": fam,
                     "major_axis": ma, "n_rows": nr, "n_cols": nc, "ro_s": ro, "co_s": co, "ctr": ctr, "key": key, "next": nxt})
    # fill_dense_submat_impl with explicit ptr / lda (test_denseskop.cc:162-296 style windows)
    k = 0
    for fam in "GU":
        for (ncols, nsr, nsc, ptr_, lda) in [(12, 4, 5, 13, 0), (13, 3, 13, 0, 0), (13, 3, 6, 29, 8), (10, 6, 3, 47, 0), (7, 5, 7, 7, 9),
                                              (4, 9, 4, 8, 0), (5, 2, 2, 3, 0), (100, 7, 33, 250, 40)]:
            ctr, key = [3, 0, 0, 0], [7 + k, 1]
            rc, out, nxt = ol.fill_dense_submat("f32", fam, ncols, nsr, nsc, ptr_, ctr, key, lda, which="ref")
            assert rc == 0
            z[f"sm{k}"] = bits(out)
            meta.append({"kind": "submat", "id": f"sm{k}", "family": fam, "n_cols": ncols, "n_srows": nsr, "n_scols": nsc, "ptr": ptr_,
                         "lda": lda, "ctr": ctr, "key": key, "next": nxt})
            k += 1
    # fill_sparse: test_sparseskop.cc shapes (7x20, 15x7; vec_nnz 1,2,3,7; keys 42,0,1) + a larger one
    k = 0
    for (r, c) in [(7, 20), (15, 7), (20, 7), (7, 15), (12, 12)]:
        for ma in "SL":
            for vn in (1, 2, 3, 7):
                for key0 in (42, 0, 1):
                    for (dt, it) in (("f64", "i64"), ("f32", "i32")):
                        if (dt, it) == ("f32", "i32") and key0 != 42:
                            continue
                        rc, rows, cols, vals, nxt = ol.fill_sparse(dt, it, r, c, vn, ma, [0, 0, 0, 0], [key0, 0], which="ref")
                        assert rc == 0
                        z[f"sp{k}_rows"], z[f"sp{k}_cols"], z[f"sp{k}_vals"] = rows, cols, vals
                        meta.append({"kind": "fill_sparse", "id": f"sp{k}", "dt": dt, "it": it, "n_rows": r, "n_cols": c, "vec_nnz": vn,
                                     "major_axis": ma, "ctr": [0, 0, 0, 0], "key": [key0, 0], "next": nxt})
                        k += 1
    rc, rows, cols, vals, nxt = ol.fill_sparse("f64", "i64", 64, 3000, 8, "S", [5, 0, 0, 0], [1997, 0], which="ref")
    z["spbig_rows"], z["spbig_cols"], z["spbig_vals"] = rows, cols, vals
    meta.append({"kind": "fill_sparse", "id": "spbig", "dt": "f64", "it": "i64", "n_rows": 64, "n_cols": 3000, "vec_nnz": 8,
                 "major_axis": "S", "ctr": [5, 0, 0, 0], "key": [1997, 0], "next": nxt})
    np.savez_compressed(os.path.join(HERE, "ref_fills.npz"), **z)
    json.dump(meta, open(os.path.join(HERE, "ref_fills.json"), "w"))


if __name__ == "__main__":
    kat()
    fills()
    print("golden fixtures written")
