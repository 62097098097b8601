"""Host-side mirror of the RandBLAS interface for the sketching hot path, over the rbx C ABI.

Same names, argument order and error behaviour as the reference headers (RandBLAS/base.hh,
dense_skops.hh, sparse_skops.hh, skge.hh, skve.hh, sksy.hh, sparse_data/sksp.hh) so that the parity
tests read like the reference's own. Buffers are torch tensors: CUDA tensors go straight to the
device entry points; CPU tensors go through the rbx_host_* entry points, which stage through HBM.
All arithmetic happens in librbx.so on the GPU -- there is no Python or CPU implementation here.
"""
import ctypes
import enum

import torch

from . import _capi
from ._capi import lib, check


class Error(RuntimeError):
    """RandBLAS::exceptions::Error (exceptions.hh:45-70)."""


class Layout(str, enum.Enum):
    ColMajor = "C"
    RowMajor = "R"


class Op(str, enum.Enum):
    NoTrans = "N"
    Trans = "T"


class MajorAxis(str, enum.Enum):      # base.hh:138-150
    Short = "S"
    Long = "L"


class DenseDistName(str, enum.Enum):  # dense_skops.hh:204-230
    Gaussian = "G"
    Uniform = "U"
    BlackBox = "B"


def _c(ch):
    return (ch.value if isinstance(ch, enum.Enum) else ch).encode()


def _u32(vals):
    return (ctypes.c_uint32 * len(vals))(*[int(v) & 0xFFFFFFFF for v in vals])


def _ptr(t):
    return ctypes.c_void_p(0) if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _call(name, *args):
    rc = getattr(lib, name)(*args)
    if rc == -1:
        raise Error(f"{name}: a condition the reference requires did not hold")
    check(rc, name)


def _suf(dtype):
    if dtype == torch.float32:
        return "f32"
    if dtype == torch.float64:
        return "f64"
    raise TypeError(f"unsupported scalar type {dtype}")


def _isuf(dtype):
    if dtype == torch.int32:
        return "i32"
    if dtype == torch.int64:
        return "i64"
    raise TypeError(f"unsupported index type {dtype}")


def _scalar(dtype, v):
    return ctypes.c_float(v) if dtype == torch.float32 else ctypes.c_double(v)


class RNGState:
    """RNGState<r123::Philox4x32> (base.hh:161-232): 128-bit counter + 64-bit key."""

    def __init__(self, key=0, counter=None):
        if isinstance(key, RNGState):
            self.counter, self.key = list(key.counter), list(key.key)
            return
        self.counter = [0, 0, 0, 0] if counter is None else [int(c) & 0xFFFFFFFF for c in counter]
        self.key = [int(key) & 0xFFFFFFFF, 0] if not isinstance(key, (list, tuple)) else [int(k) & 0xFFFFFFFF for k in key]

    def copy(self):
        return RNGState(self)

    def incr(self, n):
        c = _u32(self.counter)
        lib.rbx_ctr_add(c, ctypes.c_uint64(n))
        out = self.copy()
        out.counter = list(c)
        return out

    def __eq__(self, other):
        return self.counter == other.counter and self.key == other.key

    def __repr__(self):
        return f"RNGState(counter={self.counter}, key={self.key})"


class DenseDist:
    """dense_skops.hh:232-294."""

    def __init__(self, n_rows, n_cols, family=DenseDistName.Gaussian, major_axis=MajorAxis.Long):
        self.n_rows, self.n_cols = int(n_rows), int(n_cols)
        self.family, self.major_axis = DenseDistName(family), MajorAxis(major_axis)


def dist_to_layout(D):
    return Layout(lib.rbx_dist_to_layout(D.n_rows, D.n_cols, _c(D.major_axis)).decode())


def major_axis_length(D):
    return lib.rbx_major_axis_length(D.n_rows, D.n_cols, _c(D.major_axis))


def isometry_scale_factor(D):
    if isinstance(D, DenseDist):
        return float(min(D.n_rows, D.n_cols)) ** -0.5          # dense_skops.hh:318-326
    return float(D.vec_nnz) ** -0.5 if D.major_axis == MajorAxis.Short else (           # sparse_skops.hh:167-177
        (max(D.n_rows, D.n_cols) / min(D.n_rows, D.n_cols)) ** 0.5 * float(D.vec_nnz) ** -0.5)


class DenseSkOp:
    """dense_skops.hh:332-419. buff is None until fill_dense(S); sketches never need it."""

    def __init__(self, dist, state, dtype=torch.float32):
        if isinstance(state, int):
            state = RNGState(state)
        self.dist, self.dtype = dist, dtype
        self.n_rows, self.n_cols = dist.n_rows, dist.n_cols
        self.seed_state = state.copy()
        nx = _u32([0] * 4)
        if dist.family == DenseDistName.BlackBox:
            nx = _u32(state.counter)
        else:
            lib.rbx_dense_next_state(dist.n_rows, dist.n_cols, _c(dist.major_axis), _u32(state.counter), nx)
        self.next_state = RNGState(state.key, list(nx))
        self.buff = None
        self.layout = dist_to_layout(dist)


class SparseDist:
    """sparse_skops.hh:134-165."""

    def __init__(self, n_rows, n_cols, vec_nnz, major_axis=MajorAxis.Short):
        self.n_rows, self.n_cols, self.vec_nnz = int(n_rows), int(n_cols), int(vec_nnz)
        self.major_axis = MajorAxis(major_axis)


def nnz(S):
    D = S.dist
    return D.vec_nnz * (max(D.n_rows, D.n_cols) if D.major_axis == MajorAxis.Short else min(D.n_rows, D.n_cols))


class SparseSkOp:
    """sparse_skops.hh:183-377. rows/cols/vals live on the device once filled."""

    def __init__(self, dist, state, dtype=torch.float64, index_dtype=torch.int64):
        if isinstance(state, int):
            state = RNGState(state)
        self.dist, self.dtype, self.index_dtype = dist, dtype, index_dtype
        self.n_rows, self.n_cols = dist.n_rows, dist.n_cols
        self.seed_state = state.copy()
        nx = _u32([0] * 4)
        lib.rbx_sparse_next_state(dist.n_rows, dist.n_cols, dist.vec_nnz, _c(dist.major_axis), _u32(state.counter), nx)
        self.next_state = RNGState(state.key, list(nx))
        self.rows = self.cols = self.vals = None
        self.known_filled = False


# ---- sparse data containers (sparse_data/{coo,csr,csc}_matrix.hh): the argument types of sketch_sparse ----
class COOMatrix:
    def __init__(self, n_rows, n_cols, vals, rows, cols):
        self.n_rows, self.n_cols, self.nnz = int(n_rows), int(n_cols), int(vals.numel())
        self.vals, self.rows, self.cols = vals, rows.to(torch.int64), cols.to(torch.int64)
    fmt = "O"
    def _idx(self): return self.rows, self.cols


class CSRMatrix:
    def __init__(self, n_rows, n_cols, vals, rowptr, colidxs):
        self.n_rows, self.n_cols, self.nnz = int(n_rows), int(n_cols), int(vals.numel())
        self.vals, self.rowptr, self.colidxs = vals, rowptr.to(torch.int64), colidxs.to(torch.int64)
    fmt = "R"
    def _idx(self): return self.rowptr, self.colidxs


class CSCMatrix:
    def __init__(self, n_rows, n_cols, vals, rowidxs, colptr):
        self.n_rows, self.n_cols, self.nnz = int(n_rows), int(n_cols), int(vals.numel())
        self.vals, self.rowidxs, self.colptr = vals, rowidxs.to(torch.int64), colptr.to(torch.int64)
    fmt = "C"
    def _idx(self): return self.rowidxs, self.colptr


def _regroup(A, by):
    """device-side conversion to compressed columns ('C') or rows ('R'): rbx_sparse_regroup_* (conversions.hh:43-103)"""
    n_outer = A.n_cols if by == "C" else A.n_rows
    i0, i1 = A._idx()
    ptr = torch.empty(n_outer + 1, dtype=torch.int64, device=A.vals.device)
    idx = torch.empty(A.nnz, dtype=torch.int64, device=A.vals.device)
    vals = torch.empty_like(A.vals)
    _call(f"rbx_sparse_regroup_{_suf(A.vals.dtype)}", _c(by), _c(A.fmt), A.n_rows, A.n_cols, A.nnz, _ptr(A.vals), _ptr(i0), _ptr(i1),
          _ptr(ptr), _ptr(idx), _ptr(vals), _stream(A.vals))
    return ptr, idx, vals


def coo_to_csc(coo):
    """conversions::coo_to_csc (conversions.hh:43-62): the sort and the compression run on the device"""
    ptr, idx, vals = _regroup(coo, "C")
    return CSCMatrix(coo.n_rows, coo.n_cols, vals, idx, ptr)


def coo_to_csr(coo):
    """conversions::coo_to_csr (conversions.hh:84-103)"""
    ptr, idx, vals = _regroup(coo, "R")
    return CSRMatrix(coo.n_rows, coo.n_cols, vals, ptr, idx)


def csr_to_csc(csr):
    """CSR -> CSC of the same matrix (a data-moving transpose; transpose_as_csc, conversions.hh:143-159, is the zero-copy view)"""
    ptr, idx, vals = _regroup(csr, "C")
    return CSCMatrix(csr.n_rows, csr.n_cols, vals, idx, ptr)


def csc_to_csr(csc):
    ptr, idx, vals = _regroup(csc, "R")
    return CSRMatrix(csc.n_rows, csc.n_cols, vals, ptr, idx)


def transpose_as_csr(csc):
    """conversions::transpose_as_csr (conversions.hh:125-141): the CSC arrays of A are the CSR arrays of A^T (shared memory)"""
    return CSRMatrix(csc.n_cols, csc.n_rows, csc.vals, csc.colptr, csc.rowidxs)


def transpose_as_csc(csr):
    """conversions::transpose_as_csc (conversions.hh:143-159)"""
    return CSCMatrix(csr.n_cols, csr.n_rows, csr.vals, csr.colidxs, csr.rowptr)


# ---- fills ---------------------------------------------------------------------------------------------
def fill_dense_submat(family, n_cols, smat, n_srows, n_scols, ptr, seed, lda=0):
    """dense::fill_dense_submat_impl (dense_skops.hh:96-170); returns the next state."""
    nx = _u32([0] * 4)
    _call(f"rbx_fill_dense_submat_{_suf(smat.dtype)}", _c(family), n_cols, _ptr(smat), n_srows, n_scols, ptr,
          _u32(seed.counter), _u32(seed.key), lda, nx, _stream(smat))
    return RNGState(seed.key, list(nx))


def fill_dense(*args, device=None):
    """fill_dense(layout, D, n_rows, n_cols, ro_s, co_s, buff, seed) | fill_dense(D, buff, seed) | fill_dense(S)
    (dense_skops.hh:486-532, 565-568, 583-592)."""
    if len(args) == 1:
        S = args[0]
        if S.dist.family == DenseDistName.BlackBox:
            raise ValueError("fill_dense cannot be called with the BlackBox distribution.")
        if S.buff is not None:
            raise Error("S.buff must be empty")                      # randblas_require(!S.buff), dense_skops.hh:585
        dev = device or torch.device("cuda", torch.cuda.current_device())
        S.buff = torch.empty(S.n_rows * S.n_cols, dtype=S.dtype, device=dev)
        fill_dense(S.dist, S.buff, S.seed_state)
        return S.next_state
    if len(args) == 3:
        D, buff, seed = args
        return fill_dense(dist_to_layout(D), D, D.n_rows, D.n_cols, 0, 0, buff, seed)
    layout, D, n_rows, n_cols, ro_s, co_s, buff, seed = args
    if D.family == DenseDistName.BlackBox:
        raise ValueError("fill_dense cannot be called with the BlackBox distribution.")
    nx = _u32([0] * 4)
    common = (_c(layout), D.n_rows, D.n_cols, _c(D.family), _c(D.major_axis), n_rows, n_cols, ro_s, co_s, _ptr(buff),
              _u32(seed.counter), _u32(seed.key), nx)
    if buff.is_cuda:
        _call(f"rbx_fill_dense_{_suf(buff.dtype)}", *common, _stream(buff))
    else:
        _call(f"rbx_host_fill_dense_{_suf(buff.dtype)}", *common)
    return RNGState(seed.key, list(nx))


def repeated_fisher_yates(state, vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor=None, vals=None):
    """sparse::repeated_fisher_yates (sparse_skops.hh:53-113); returns its input state like the reference (:105)."""
    vdt = vals.dtype if vals is not None else torch.float64
    _call(f"rbx_repeated_fisher_yates_{_suf(vdt)}_{_isuf(idxs_major.dtype)}", _u32(state.counter), _u32(state.key),
          vec_nnz, dim_major, dim_minor, _ptr(idxs_major), _ptr(idxs_minor), _ptr(vals), _stream(idxs_major))
    return state.copy()


def sample_indices_iid_uniform(n, k, samples, state):
    """util::sample_indices_iid_uniform (util.hh:311-334): k iid samples of {0..n-1} into the int64 device tensor
    `samples`; returns the next state."""
    nx = _u32([0] * 4)
    _call("rbx_sample_indices_iid_uniform", n, k, _ptr(samples), _u32(state.counter), _u32(state.key), nx, _stream(samples))
    return RNGState(state.key, list(nx))


def sample_indices_iid(n, cdf, k, samples, state):
    """util::sample_indices_iid (util.hh:284-305): k iid samples from the distribution whose cdf is the device tensor `cdf`."""
    nx = _u32([0] * 4)
    _call(f"rbx_sample_indices_iid_{_suf(cdf.dtype)}", n, _ptr(cdf), k, _ptr(samples), _u32(state.counter), _u32(state.key), nx,
          _stream(samples))
    return RNGState(state.key, list(nx))


def weights_to_cdf(n, w, error_if_below=None):
    """util::weights_to_cdf (util.hh:257-270), in place on the device tensor w; raises Error where the reference throws."""
    if error_if_below is None:
        error_if_below = -float(torch.finfo(w.dtype).eps)
    _call(f"rbx_weights_to_cdf_{_suf(w.dtype)}", n, _ptr(w), error_if_below, _stream(w))


def fill_sparse(S, device=None):
    """fill_sparse(SparseSkOp&) (sparse_skops.hh:389-413)."""
    dev = device or torch.device("cuda", torch.cuda.current_device())
    n = nnz(S)
    if S.rows is None:
        S.rows = torch.empty(n, dtype=S.index_dtype, device=dev)
        S.cols = torch.empty(n, dtype=S.index_dtype, device=dev)
        S.vals = torch.empty(n, dtype=S.dtype, device=dev)
    D = S.dist
    _call(f"rbx_fill_sparse_{_suf(S.dtype)}_{_isuf(S.index_dtype)}", D.n_rows, D.n_cols, D.vec_nnz, _c(D.major_axis),
          _u32(S.seed_state.counter), _u32(S.seed_state.key), _ptr(S.rows), _ptr(S.cols), _ptr(S.vals),
          ctypes.c_void_p(0), _stream(S.rows))
    S.known_filled = True


def coo_view_of_skop(S):
    """sparse_skops.hh:483-490."""
    if not S.known_filled:
        raise Error("S.known_filled was required")
    return COOMatrix(S.n_rows, S.n_cols, S.vals, S.rows, S.cols)


# ---- sketches --------------------------------------------------------------------------------------------
def _is_skop(x):
    return isinstance(x, (DenseSkOp, SparseSkOp))


def _sketch(side, layout, opS, opA, p0, p1, p2, alpha, S, ro_s, co_s, A, lda, beta, B, ldb):
    dt = B.dtype
    suf = _suf(dt)
    head = (_c(side), _c(layout), _c(opS), _c(opA), p0, p1, p2, _scalar(dt, alpha))
    tail = (ro_s, co_s, _ptr(A), lda, _scalar(dt, beta), _ptr(B), ldb)
    host = not B.is_cuda
    if isinstance(S, DenseSkOp):
        D = S.dist
        if D.family == DenseDistName.BlackBox and S.buff is None:
            raise ValueError("a BlackBox operator needs an explicit buffer")
        fam = D.family if D.family != DenseDistName.BlackBox else DenseDistName.Gaussian
        op = (D.n_rows, D.n_cols, _c(fam), _c(D.major_axis), _u32(S.seed_state.counter), _u32(S.seed_state.key))
        if host:
            if S.buff is not None:
                raise Error("host-pointer sketches take implicit operators only")
            _call(f"rbx_host_sketch_dense_{suf}", *head, *op, *tail)
        else:
            _call(f"rbx_sketch_dense_{suf}", *head, *op, _ptr(S.buff), _c(S.layout), *tail, _stream(B))
    elif isinstance(S, SparseSkOp):
        D = S.dist
        if S.known_filled and host:
            # explicit (possibly caller-edited) triplets with host data: stage A and B and take the same COO path as device data
            dev = S.vals.device
            rowsA = (p2 if Op(opA) == Op.NoTrans else p1) if side == "L" else (p0 if Op(opA) == Op.NoTrans else p2)
            Ad = A.to(dev, non_blocking=True)
            Bd = B.to(dev, non_blocking=True)
            _sketch(side, layout, opS, opA, p0, p1, p2, alpha, S, ro_s, co_s, Ad, lda, beta, Bd, ldb)
            B.copy_(Bd)
            return
        if S.known_filled and not host:
            # explicit triplets (possibly modified by the caller): left_spmm / right_spmm on the COO view
            r64, c64 = S.rows.to(torch.int64), S.cols.to(torch.int64)
            _call(f"rbx_spmm_coo_{suf}", *head, D.n_rows, D.n_cols, nnz(S), _ptr(r64), _ptr(c64), _ptr(S.vals), *tail, _stream(B))
            torch.cuda.current_stream(B.device).synchronize() if r64 is not S.rows else None
        else:
            op = (D.n_rows, D.n_cols, D.vec_nnz, _c(D.major_axis), _u32(S.seed_state.counter), _u32(S.seed_state.key))
            if host:
                _call(f"rbx_host_sketch_sparseop_{suf}", *head, *op, *tail)
            else:
                _call(f"rbx_sketch_sparseop_{suf}", *head, *op, *tail, _stream(B))
    else:
        raise TypeError("S must be a DenseSkOp or SparseSkOp")


def sketch_general(layout, op1, op2, p0, p1, p2, alpha, *rest):
    """The six overloads of skge.hh:771-1214.
    left : sketch_general(layout, opS, opA, d, n, m, alpha, S, [ro_s, co_s,] A, lda, beta, B, ldb)
    right: sketch_general(layout, opA, opS, m, d, n, alpha, A, lda, S, [ro_s, co_s,] beta, B, ldb)"""
    if _is_skop(rest[0]):
        opS, opA, d, n, m = op1, op2, p0, p1, p2
        if len(rest) == 8:
            S, ro_s, co_s, A, lda, beta, B, ldb = rest
        else:
            S, A, lda, beta, B, ldb = rest
            ro_s = co_s = 0
            rows, cols = (d, m) if Op(opS) == Op.NoTrans else (m, d)
            if S.dist.n_rows != rows or S.dist.n_cols != cols:      # skge.hh:1104-1111
                raise Error("S.dist dims must match (d, m)")
        return _sketch("L", layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, lda, beta, B, ldb)
    opA, opS, m, d, n = op1, op2, p0, p1, p2
    if len(rest) == 8:
        A, lda, S, ro_s, co_s, beta, B, ldb = rest
    else:
        A, lda, S, beta, B, ldb = rest
        ro_s = co_s = 0
        rows, cols = (n, d) if Op(opS) == Op.NoTrans else (d, n)
        if S.dist.n_rows != rows or S.dist.n_cols != cols:          # skge.hh:1206-1213
            raise Error("S.dist dims must match (n, d)")
    return _sketch("R", layout, opS, opA, m, d, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb)


def sketch_vector(opS, *rest):
    """skve.hh:152-176 (submatrix form: opS, d, m, alpha, S, ro_s, co_s, x, incx, beta, y, incy) and
    skve.hh:244-258 (full form: opS, alpha, S, x, incx, beta, y, incy)."""
    if len(rest) == 11:
        d, m, alpha, S, ro_s, co_s, x, incx, beta, y, incy = rest
    else:
        alpha, S, x, incx, beta, y, incy = rest
        d, m, ro_s, co_s = S.dist.n_rows, S.dist.n_cols, 0, 0
    _d, _m = (m, d) if Op(opS) == Op.Trans else (d, m)
    return sketch_general(Layout.RowMajor, opS, Op.NoTrans, _d, 1, _m, alpha, S, ro_s, co_s, x, incx, beta, y, incy)


def require_symmetric(layout, A, n, lda, tol):
    """util::require_symmetric (util.hh:165-188): a precondition check, not part of the timed path. Device tensors are
    checked by one kernel over the upper triangle (rbx_require_symmetric_*); host tensors on the host, like the reference."""
    if tol < 0:
        return
    if A.is_cuda:
        rc = getattr(lib, f"rbx_require_symmetric_{_suf(A.dtype)}")(_c(layout), _ptr(A), n, lda, _scalar(A.dtype, tol), _stream(A))
        if rc == -1:
            raise Error("Symmetry check failed.")
        check(rc, "rbx_require_symmetric")
        return
    M = torch.as_strided(A, (n, n), (1, lda) if Layout(layout) == Layout.ColMajor else (lda, 1))
    viol = (M - M.T).abs()
    rel = (M.abs() + M.T.abs() + 1) * tol
    if bool((viol > rel).any()):
        raise Error("Symmetry check failed.")


def sketch_symmetric(layout, *rest, sym_check_tol=0):
    """sksy.hh:165-184 / 300-319 (submatrix forms) and 413-430 / 520-537 (full forms).
    right: (layout, n, d, alpha, A, lda, S, [ro_s, co_s,] beta, B, ldb); left: (layout, d, n, alpha, S, [ro_s, co_s,] A, lda, beta, B, ldb)"""
    idx = [i for i, r in enumerate(rest) if _is_skop(r)][0]
    if idx == 1 and len(rest) == 7:        # full left form: (layout, alpha, S, A, lda, beta, B, ldb)
        alpha, S, A, lda, beta, B, ldb = rest
        d, n, ro_s, co_s, left = S.dist.n_rows, S.dist.n_cols, 0, 0, True
    elif idx == 3 and len(rest) == 11:     # left submatrix form
        d, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb = rest
        left = True
    elif idx == 3 and len(rest) == 7:      # full right form: (layout, alpha, A, lda, S, beta, B, ldb)
        alpha, A, lda, S, beta, B, ldb = rest
        n, d, ro_s, co_s, left = S.dist.n_rows, S.dist.n_cols, 0, 0, False
    elif idx == 5 and len(rest) == 11:     # right submatrix form
        n, d, alpha, A, lda, S, ro_s, co_s, beta, B, ldb = rest
        left = False
    else:
        raise TypeError("bad sketch_symmetric arguments")
    require_symmetric(layout, A, n, lda, sym_check_tol)
    if left:
        return sketch_general(layout, Op.NoTrans, Op.NoTrans, d, n, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb)
    return sketch_general(layout, Op.NoTrans, Op.NoTrans, n, d, n, alpha, A, lda, S, ro_s, co_s, beta, B, ldb)


def _coo_triplets(A):
    """(rows, cols, vals) of a COO / CSR / CSC container as int64 device arrays (rbx_expand_ptr expands a pointer array)"""
    i0, i1 = A._idx()
    if A.fmt == "O":
        return i0, i1, A.vals
    outer = torch.empty(max(A.nnz, 1), dtype=torch.int64, device=A.vals.device)
    n_outer = A.n_rows if A.fmt == "R" else A.n_cols
    _call("rbx_expand_ptr", n_outer, _ptr(i0 if A.fmt == "R" else i1), _ptr(outer), _stream(A.vals))
    return (outer[: A.nnz], i1, A.vals) if A.fmt == "R" else (i0, outer[: A.nnz], A.vals)


def left_spmm(layout, opA, opB, d, n, m, alpha, A, ro_a, co_a, B, ldb, beta, C, ldc):
    """sparse_data::left_spmm (spmm_dispatch.hh:48-160): C (d x n) = alpha * op(submat(A)) (d x m, sparse) * op(B) (m x n) + beta * C
    for a COOMatrix / CSRMatrix / CSCMatrix A on the device. Output-stationary gather, no atomics (rbx_spmm_coo_*)."""
    if A.fmt != "O" and (ro_a != 0 or co_a != 0):
        raise Error("CSR / CSC operands are taken whole (spmm_dispatch.hh:97-104)")
    dt = C.dtype
    if A.fmt != "O":
        # compressed operands go down as they are: used in place when compressed along the output axis (rbx_spmm_csx_*)
        ptr, idx = (A.rowptr, A.colidxs) if A.fmt == "R" else (A.colptr, A.rowidxs)
        _call(f"rbx_spmm_csx_{_suf(dt)}", b"L", _c(layout), _c(opA), _c(opB), d, n, m, _scalar(dt, alpha), A.n_rows, A.n_cols, A.nnz,
              A.fmt.encode(), _ptr(ptr), _ptr(idx), _ptr(A.vals), _ptr(B), ldb, _scalar(dt, beta), _ptr(C), ldc, _stream(C))
        return
    rows, cols, vals = _coo_triplets(A)
    _call(f"rbx_spmm_coo_{_suf(dt)}", b"L", _c(layout), _c(opA), _c(opB), d, n, m, _scalar(dt, alpha), A.n_rows, A.n_cols, A.nnz,
          _ptr(rows), _ptr(cols), _ptr(vals), ro_a, co_a, _ptr(B), ldb, _scalar(dt, beta), _ptr(C), ldc, _stream(C))


def right_spmm(layout, opA, opB, m, d, n, alpha, A, lda, B, ro_b, co_b, beta, C, ldc):
    """sparse_data::right_spmm (spmm_dispatch.hh:162-200): C (m x d) = alpha * op(A) (m x n, dense) * op(submat(B)) (n x d, sparse)
    + beta * C, as the transposed left product in the opposite layout."""
    t_opB = Op.Trans if Op(opB) == Op.NoTrans else Op.NoTrans
    t_layout = Layout.RowMajor if Layout(layout) == Layout.ColMajor else Layout.ColMajor
    left_spmm(t_layout, t_opB, opA, d, m, n, alpha, B, ro_b, co_b, A, lda, beta, C, ldc)


def sketch_sparse(layout, op1, op2, p0, p1, p2, alpha, *rest):
    """sksp.hh:464-485 (left: layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb) and
    sksp.hh:595-616 (right: layout, opA, opS, m, d, n, alpha, A, ro_a, co_a, S, ro_s, co_s, beta, B, ldb)."""
    if _is_skop(rest[0]):
        side, opS, opA = "L", op1, op2
        S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb = rest
    else:
        side, opA, opS = "R", op1, op2
        A, ro_a, co_a, S, ro_s, co_s, beta, B, ldb = rest
    D = S.dist
    dt = B.dtype
    if S.buff is not None or D.family == DenseDistName.BlackBox:
        # the reference multiplies by S.buff here (sksp.hh:166-172); the device path regenerates the operator from its seed,
        # which an explicit buffer may not match
        raise Error("sketch_sparse takes an implicit DenseSkOp (S.buff is None) on the device; use sketch_general on A^T for an explicit operator")
    if not B.is_cuda:
        # host arrays (what a caller of the reference owns): rbx_host_sketch_sparsedata_* stages the sparse matrix and B
        # through HBM around the device call
        i0, i1 = A._idx()
        _call(f"rbx_host_sketch_sparsedata_{_suf(dt)}", _c(side), _c(layout), _c(opS), _c(opA), p0, p1, p2, _scalar(dt, alpha),
              D.n_rows, D.n_cols, _c(D.family), _c(D.major_axis), _u32(S.seed_state.counter), _u32(S.seed_state.key),
              ro_s, co_s, A.fmt.encode(), A.n_rows, A.n_cols, A.nnz, _ptr(A.vals), _ptr(i0), _ptr(i1), ro_a, co_a,
              _scalar(dt, beta), _ptr(B), ldb)
        return
    i0, i1 = A._idx()
    _call(f"rbx_sketch_sparsedata_{_suf(dt)}", _c(side), _c(layout), _c(opS), _c(opA), p0, p1, p2, _scalar(dt, alpha),
          D.n_rows, D.n_cols, _c(D.family), _c(D.major_axis), _u32(S.seed_state.counter), _u32(S.seed_state.key),
          ro_s, co_s, A.fmt.encode(), A.n_rows, A.n_cols, A.nnz, _ptr(A.vals), _ptr(i0), _ptr(i1), ro_a, co_a,
          _scalar(dt, beta), _ptr(B), ldb, _stream(B))
