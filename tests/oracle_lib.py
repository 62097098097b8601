"""Test-side loader for the CPU checkers under oracle/ (TEST INFRASTRUCTURE ONLY).

  ORC : oracle/_build/liboracle.so      -- the plain-C restatement (oracle/randblas_oracle.c)
  REF : oracle/_ref/librandblas_ref.so  -- the unmodified reference headers compiled against oracle/shims
                                           (None when it has not been built)
Nothing in randblas_b200/ imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_PATH = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "librandblas_ref.so")

if not os.path.exists(ORC_PATH):
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True, capture_output=True)
ORC = ctypes.CDLL(ORC_PATH)
REF = ctypes.CDLL(REF_PATH) if os.path.exists(REF_PATH) else None

c_i64, c_u64, c_char, c_vp = ctypes.c_int64, ctypes.c_uint64, ctypes.c_char, ctypes.c_void_p
ORC.orc_uneg11.restype = ctypes.c_float
ORC.orc_uneg11.argtypes = [ctypes.c_uint32]
ORC.orc_u01.restype = ctypes.c_float
ORC.orc_u01.argtypes = [ctypes.c_uint32]
for _n in ("orc_sweep_sincos", "orc_sweep_log"):
    getattr(ORC, _n).restype = c_i64
    getattr(ORC, _n).argtypes = [c_u64, c_u64, c_u64, c_vp]


def u32(vals):
    return (ctypes.c_uint32 * len(vals))(*[int(v) & 0xFFFFFFFF for v in vals])


def ptr(a):
    return c_vp(0) if a is None else a.ctypes.data_as(c_vp)


def ch(c):
    return c_char(c.encode() if isinstance(c, str) else c)


def _np(dt):
    return {"f32": np.float32, "f64": np.float64}[dt]


def _sc(dt, v):
    return ctypes.c_float(v) if dt == "f32" else ctypes.c_double(v)


def _lib(which):
    if which == "ref":
        if REF is None:
            raise RuntimeError("oracle/_ref/librandblas_ref.so not built")
        return REF, "ref_"
    return ORC, "orc_"


def philox(ctr, key, which="orc"):
    out = u32([0] * 4)
    if which == "ref":
        REF.ref_philox4x32(u32(ctr), u32(key), out)
    else:
        ORC.orc_philox4x32_10(u32(ctr), u32(key), out)
    return list(out)


def ctr_add(ctr, n, which="orc"):
    c = u32(ctr)
    if which == "ref":
        REF.ref_ctr_incr(c, c_u64(n))
    else:
        ORC.orc_ctr_add(c, c_u64(n))
    return list(c)


def generate4(family, ctr, key, which="orc"):
    out = (ctypes.c_float * 4)()
    lib, pre = _lib(which)
    getattr(lib, pre + "generate4")(ch(family), u32(ctr), u32(key), out)
    return np.array(list(out), dtype=np.float32)


def boxmuller_model(u0, u1):
    """the explicit-FMA models of glibc sincosf/logf in the oracle (what the CUDA kernel implements)"""
    s, c = ctypes.c_float(), ctypes.c_float()
    ORC.orc_boxmuller_model(ctypes.c_uint32(u0), ctypes.c_uint32(u1), ctypes.byref(s), ctypes.byref(c))
    return s.value, c.value


def boxmuller_arr(u0_arr, u1_arr, model=False):
    """r123::boxmuller on arrays: live host libm (model=False) or the explicit-FMA models (model=True)"""
    u0 = np.ascontiguousarray(u0_arr, dtype=np.uint32)
    u1 = np.ascontiguousarray(u1_arr, dtype=np.uint32)
    s = np.empty(len(u0), np.float32)
    c = np.empty(len(u0), np.float32)
    ORC.orc_boxmuller_arr(c_i64(len(u0)), ptr(u0), ptr(u1), ptr(s), ptr(c), ctypes.c_int(1 if model else 0))
    return s, c


def fill_dense_submat(dt, family, n_cols, n_srows, n_scols, ptr_, ctr, key, lda=0, which="orc"):
    lib, pre = _lib(which)
    ld = lda if lda > 0 else n_scols
    out = np.full(max(n_srows * ld, 1), np.nan, dtype=_np(dt))
    nxt = u32([0] * 4)
    rc = getattr(lib, f"{pre}fill_dense_submat_{dt}")(ch(family), c_i64(n_cols), ptr(out), c_i64(n_srows), c_i64(n_scols), c_i64(ptr_),
                                                     u32(ctr), u32(key), c_i64(lda), nxt)
    return rc, out, list(nxt)


def fill_dense(dt, layout, D_rows, D_cols, family, major_axis, n_rows, n_cols, ro_s, co_s, ctr, key, which="orc"):
    lib, pre = _lib(which)
    out = np.full(max(n_rows * n_cols, 1), np.nan, dtype=_np(dt))
    nxt = u32([0] * 4)
    rc = getattr(lib, f"{pre}fill_dense_{dt}")(ch(layout), c_i64(D_rows), c_i64(D_cols), ch(family), ch(major_axis), c_i64(n_rows),
                                              c_i64(n_cols), c_i64(ro_s), c_i64(co_s), ptr(out), u32(ctr), u32(key), nxt)
    return rc, out[: n_rows * n_cols], list(nxt)


def fill_sparse(dt, it, n_rows, n_cols, vec_nnz, major_axis, ctr, key, which="orc"):
    lib, pre = _lib(which)
    lng, sht = max(n_rows, n_cols), min(n_rows, n_cols)
    nnz = vec_nnz * (lng if major_axis == "S" else sht)
    idt = np.int32 if it == "i32" else np.int64
    rows, cols, vals = np.full(nnz, -1, idt), np.full(nnz, -1, idt), np.zeros(nnz, _np(dt))
    nxt = u32([0] * 4)
    rc = getattr(lib, f"{pre}fill_sparse_{dt}_{it}")(c_i64(n_rows), c_i64(n_cols), c_i64(vec_nnz), ch(major_axis), u32(ctr), u32(key),
                                                    ptr(rows), ptr(cols), ptr(vals), nxt)
    return rc, rows, cols, vals, list(nxt)


def sketch_dense(dt, side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                 A, lda, beta, B, ldb, which="orc"):
    """B is modified in place (numpy array)."""
    if which == "ref":
        return REF.__getattr__(f"ref_sketch_dense_{dt}")(ch(side), ch(layout), ch(opS), ch(opA), c_i64(p0), c_i64(p1), c_i64(p2), _sc(dt, alpha),
                                                        c_i64(D_rows), c_i64(D_cols), ch(family), ch(major_axis), u32(ctr), u32(key),
                                                        c_vp(0), ch("C"), c_i64(ro_s), c_i64(co_s), ptr(A), c_i64(lda), _sc(dt, beta), ptr(B), c_i64(ldb))
    return ORC.__getattr__(f"orc_sketch_dense_{dt}")(ch(side), ch(layout), ch(opS), ch(opA), c_i64(p0), c_i64(p1), c_i64(p2), _sc(dt, alpha),
                                                    c_i64(D_rows), c_i64(D_cols), ch(family), ch(major_axis), u32(ctr), u32(key),
                                                    c_i64(ro_s), c_i64(co_s), ptr(A), c_i64(lda), _sc(dt, beta), ptr(B), c_i64(ldb))


def sketch_sparseop(dt, side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, vec_nnz, major_axis, ctr, key, ro_s, co_s,
                    A, lda, beta, B, ldb, which="orc"):
    args = [ch(side), ch(layout), ch(opS), ch(opA), c_i64(p0), c_i64(p1), c_i64(p2), _sc(dt, alpha), c_i64(S_rows), c_i64(S_cols),
            c_i64(vec_nnz), ch(major_axis), u32(ctr), u32(key), c_i64(ro_s), c_i64(co_s), ptr(A), c_i64(lda), _sc(dt, beta), ptr(B), c_i64(ldb)]
    if which == "ref":
        return REF.__getattr__(f"ref_sketch_sparseop_{dt}")(*args, c_vp(0), c_vp(0))
    return ORC.__getattr__(f"orc_sketch_sparseop_{dt}")(*args)


def sketch_sparsedata(dt, side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                      fmt, A_rows, A_cols, vals, idx0, idx1, ro_a, co_a, beta, B, ldb, which="orc"):
    lib, pre = _lib(which)
    return getattr(lib, f"{pre}sketch_sparsedata_{dt}")(
        ch(side), ch(layout), ch(opS), ch(opA), c_i64(p0), c_i64(p1), c_i64(p2), _sc(dt, alpha), c_i64(D_rows), c_i64(D_cols), ch(family),
        ch(major_axis), u32(ctr), u32(key), c_i64(ro_s), c_i64(co_s), ch(fmt), c_i64(A_rows), c_i64(A_cols), c_i64(len(vals)),
        ptr(vals), ptr(idx0), ptr(idx1), c_i64(ro_a), c_i64(co_a), _sc(dt, beta), ptr(B), c_i64(ldb))


def dense_next_state(n_rows, n_cols, major_axis, ctr):
    out = u32([0] * 4)
    ORC.orc_dense_next_state(c_i64(n_rows), c_i64(n_cols), ch(major_axis), u32(ctr), out)
    return list(out)


# ---- index sampling utilities (util.hh:257-334) -----------------------------------------------------
def sample_indices_iid_uniform(n, k, ctr, key, which="orc"):
    lib, pre = _lib(which)
    out = np.zeros(max(k, 1), np.int64)
    nx = u32([0] * 4)
    rc = getattr(lib, pre + "sample_indices_iid_uniform")(ctypes.c_int64(n), ctypes.c_int64(k), ptr(out), u32(ctr), u32(key), nx)
    return rc, out[:k], list(nx)


def sample_indices_iid(dt, cdf, k, ctr, key, which="orc"):
    lib, pre = _lib(which)
    cdf = np.ascontiguousarray(cdf, _np(dt))
    out = np.zeros(max(k, 1), np.int64)
    nx = u32([0] * 4)
    rc = getattr(lib, f"{pre}sample_indices_iid_{dt}")(ctypes.c_int64(len(cdf)), ptr(cdf), ctypes.c_int64(k), ptr(out), u32(ctr), u32(key), nx)
    return rc, out[:k], list(nx)


def weights_to_cdf(dt, w, error_if_below=None, which="orc"):
    lib, pre = _lib(which)
    w = np.array(w, _np(dt))
    eib = -np.finfo(_np(dt)).eps if error_if_below is None else error_if_below
    rc = getattr(lib, f"{pre}weights_to_cdf_{dt}")(ctypes.c_int64(len(w)), ptr(w), _sc(dt, eib))
    return rc, w
