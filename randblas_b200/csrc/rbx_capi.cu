// C ABI (include/rbx.h): argument validation mirroring the reference's randblas_require checks, index
// algebra that the reference does on the host (offset_and_ldim, dims_before_op, layout flips), and
// dispatch to the sm_100a kernels. Nothing here computes sketches on the CPU.
#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include "rbx_kernels.cuh"

namespace rbx {

static std::atomic<int64_t> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static int require_device() {
    static int cached = 0;   // 0 unknown, 1 ok, -1 none
    if (cached == 0) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess) { cudaGetLastError(); n = 0; }
        int ok = -1;
        if (n > 0) {
            int dev = 0; cudaGetDevice(&dev);
            cudaDeviceProp prop;
            if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess && prop.major == 10) ok = 1;
        }
        cached = ok;
    }
    return cached == 1 ? RBX_OK : RBX_ERR_NO_DEVICE;
}
// Every entry point validates its arguments BEFORE this macro, so a "dry run" stops here with the verdict of
// the argument checks; the host-pointer wrappers use that to report RBX_ERR_ARG before touching the device.
static thread_local bool t_dry_run = false;
struct DryRun { DryRun() { t_dry_run = true; } ~DryRun() { t_dry_run = false; } };
#define RBX_NEED_DEVICE() do { if (t_dry_run) return RBX_OK; int rc__ = require_device(); if (rc__ != RBX_OK) return rc__; } while (0)
#define RBX_REQUIRE(cond) do { if (!(cond)) return RBX_ERR_ARG; } while (0)
#define RBX_TRY(expr) do { int rc__ = (expr); if (rc__ != RBX_OK) return rc__; } while (0)

static inline bool is_layout(char c) { return c == 'C' || c == 'R'; }
static inline bool is_op(char c) { return c == 'N' || c == 'T'; }

// ---- dense fills ------------------------------------------------------------------------------------
template <typename T>
static int fill_dense_submat(char family, int64_t n_cols, T* smat, int64_t n_srows, int64_t n_scols, int64_t ptr,
                             const uint32_t* ctr, const uint32_t* key, int64_t lda, uint32_t* next_ctr, void* stream) {
    RBX_REQUIRE(family == 'G' || family == 'U');
    if (lda <= 0) lda = n_scols; else RBX_REQUIRE(lda >= n_scols);          // dense_skops.hh:106-110
    RBX_REQUIRE(n_cols >= n_scols && n_cols > 0 && ptr >= 0);               // dense_skops.hh:111
    const int64_t r0 = ptr / n_cols, c0 = ptr % n_cols, W = (n_cols + 3) / 4;
    if (next_ctr) {                                                          // dense_skops.hh:166-169
        Ctr128 nx = ctr_add(ctr_add(mk_ctr(ctr), (uint64_t)(r0 * W + c0 / 4)), (uint64_t)(n_srows * W));
        put_ctr(nx, next_ctr);
    }
    RBX_NEED_DEVICE();
    return launch_fill_window<T>(family, n_cols, smat, n_srows, n_scols, r0, c0, mk_ctr(ctr), mk_key(key), lda, 1, false,
                                 (cudaStream_t) stream);
}

template <typename T>
static int fill_dense(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis, int64_t n_rows, int64_t n_cols,
                      int64_t ro_s, int64_t co_s, T* buff, const uint32_t* ctr, const uint32_t* key, uint32_t* next_ctr, void* stream) {
    RBX_REQUIRE(is_layout(layout));
    RBX_REQUIRE(family == 'G' || family == 'U');                            // BlackBox / unknown: dense_skops.hh:515-520
    RBX_REQUIRE(major_axis == 'L' || major_axis == 'S');
    RBX_REQUIRE(n_rows >= 0 && n_cols >= 0 && ro_s >= 0 && co_s >= 0);
    RBX_REQUIRE(D_rows >= n_rows + ro_s && D_cols >= n_cols + co_s);        // dense_skops.hh:488-489
    const int64_t ma = major_axis_length(D_rows, D_cols, major_axis), W = (ma + 3) / 4;
    const bool nat_row = natural_is_rowmajor(D_rows, D_cols, major_axis);
    // natural RowMajor: parent = S; natural ColMajor: parent = S^T (dense_skops.hh:494-503)
    const int64_t r0 = nat_row ? ro_s : co_s, c0 = nat_row ? co_s : ro_s;
    const int64_t nsr = nat_row ? n_rows : n_cols, nsc = nat_row ? n_cols : n_rows;
    if (next_ctr) {
        Ctr128 nx = ctr_add(ctr_add(mk_ctr(ctr), (uint64_t)(r0 * W + c0 / 4)), (uint64_t)(nsr * W));
        put_ctr(nx, next_ctr);
    }
    // strides of buff for (S row, S col), then for parent (row, col)
    const int64_t s_r = (layout == 'C') ? 1 : n_cols, s_c = (layout == 'C') ? n_rows : 1;
    const int64_t rs = nat_row ? s_r : s_c, cs = nat_row ? s_c : s_r;
    RBX_NEED_DEVICE();
    return launch_fill_window<T>(family, ma, buff, nsr, nsc, r0, c0, mk_ctr(ctr), mk_key(key), rs, cs, family == 'U',
                                 (cudaStream_t) stream);
}

template <typename T>
static int host_fill_dense(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis, int64_t n_rows, int64_t n_cols,
                           int64_t ro_s, int64_t co_s, T* buff_host, const uint32_t* ctr, const uint32_t* key, uint32_t* next_ctr) {
    // argument checks and the returned state, without touching the device
    { DryRun dry; RBX_TRY(fill_dense<T>(layout, D_rows, D_cols, family, major_axis, n_rows, n_cols, ro_s, co_s, (T*) nullptr, ctr, key, next_ctr, nullptr)); }
    RBX_NEED_DEVICE();
    if (n_rows == 0 || n_cols == 0) return RBX_OK;
    // The window is generated in chunks of its slow axis (about 256 MB each) on two streams: chunk i + 1 is generated while
    // chunk i crosses PCIe, and the device holds two chunks instead of the whole window (17 GB at C3).
    const int64_t fast = (layout == 'C') ? n_rows : n_cols, slow = (layout == 'C') ? n_cols : n_rows;
    int64_t per = ((int64_t) 256 << 20) / (fast * (int64_t) sizeof(T));
    if (per < 1) per = 1;
    if (per > slow) per = slow;
    const size_t chunk_bytes = (size_t) per * fast * sizeof(T);
    T* buf[2] = { nullptr, nullptr };
    cudaStream_t st[2] = { nullptr, nullptr };
    int rc = RBX_OK;
    const int nbuf = per < slow ? 2 : 1;
    for (int b = 0; b < nbuf && rc == RBX_OK; ++b) {
        if (ws_alloc(&buf[b], chunk_bytes, nullptr) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_ALLOC; }
        else if (cudaStreamCreateWithFlags(&st[b], cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    }
    if (rc == RBX_OK && cudaStreamSynchronize(nullptr) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    int64_t i = 0;
    for (int64_t s0 = 0; s0 < slow && rc == RBX_OK; s0 += per, ++i) {
        const int b = (int)(i % nbuf);
        const int64_t len = (slow - s0 < per) ? slow - s0 : per;
        if (layout == 'C') rc = fill_dense<T>(layout, D_rows, D_cols, family, major_axis, n_rows, len, ro_s, co_s + s0, buf[b], ctr, key, nullptr, (void*) st[b]);
        else               rc = fill_dense<T>(layout, D_rows, D_cols, family, major_axis, len, n_cols, ro_s + s0, co_s, buf[b], ctr, key, nullptr, (void*) st[b]);
        if (rc == RBX_OK && cudaMemcpyAsync(buff_host + s0 * fast, buf[b], (size_t) len * fast * sizeof(T), cudaMemcpyDeviceToHost, st[b]) != cudaSuccess) {
            cudaGetLastError(); rc = RBX_ERR_CUDA;
        }
    }
    for (int b = 0; b < nbuf; ++b) {
        if (st[b]) { if (cudaStreamSynchronize(st[b]) != cudaSuccess && rc == RBX_OK) { cudaGetLastError(); rc = RBX_ERR_CUDA; } cudaStreamDestroy(st[b]); }
        if (buf[b]) cudaFreeAsync(buf[b], nullptr);
    }
    return rc;
}

// ---- sparse fills -----------------------------------------------------------------------------------
template <typename T, typename I>
static int fill_sparse(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t* ctr, const uint32_t* key,
                       I* rows, I* cols, T* vals, uint32_t* next_ctr, void* stream) {
    RBX_REQUIRE(major_axis == 'L' || major_axis == 'S');
    RBX_REQUIRE(n_rows > 0 && n_cols > 0 && vec_nnz > 0);
    const int64_t lng = n_rows > n_cols ? n_rows : n_cols, sht = n_rows < n_cols ? n_rows : n_cols;
    const bool is_wide = (n_rows == sht);                                    // sparse_skops.hh:394
    I* short_idx = is_wide ? rows : cols;
    I* long_idx  = is_wide ? cols : rows;
    if (next_ctr) rbx_sparse_next_state(n_rows, n_cols, vec_nnz, major_axis, ctr, next_ctr);
    RBX_REQUIRE(vec_nnz <= (major_axis == 'S' ? sht : lng));                 // sparse_skops.hh:63
    RBX_NEED_DEVICE();
    if (major_axis == 'S')
        return launch_fisher_yates<T, I>(mk_ctr(ctr), mk_key(key), vec_nnz, sht, lng, short_idx, long_idx, vals, (cudaStream_t) stream);
    return launch_fisher_yates<T, I>(mk_ctr(ctr), mk_key(key), vec_nnz, lng, sht, long_idx, short_idx, vals, (cudaStream_t) stream);
}

template <typename T, typename I>
static int host_fill_sparse(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t* ctr, const uint32_t* key,
                            I* rows_h, I* cols_h, T* vals_h, uint32_t* next_ctr) {
    { DryRun dry; RBX_TRY((fill_sparse<T, I>(n_rows, n_cols, vec_nnz, major_axis, ctr, key, (I*) nullptr, (I*) nullptr, (T*) nullptr, nullptr, nullptr))); }
    RBX_NEED_DEVICE();
    const int64_t lng = n_rows > n_cols ? n_rows : n_cols, sht = n_rows < n_cols ? n_rows : n_cols;
    const size_t nnz = (size_t)(vec_nnz * (major_axis == 'S' ? lng : sht));
    I *rows = nullptr, *cols = nullptr; T* vals = nullptr;
    if (cudaMalloc(&rows, nnz * sizeof(I)) != cudaSuccess || cudaMalloc(&cols, nnz * sizeof(I)) != cudaSuccess ||
        cudaMalloc(&vals, nnz * sizeof(T)) != cudaSuccess) {
        cudaGetLastError(); cudaFree(rows); cudaFree(cols); cudaFree(vals);
        return RBX_ERR_ALLOC;
    }
    int rc = fill_sparse<T, I>(n_rows, n_cols, vec_nnz, major_axis, ctr, key, rows, cols, vals, next_ctr, nullptr);
    if (rc == RBX_OK) {
        if (cudaMemcpy(rows_h, rows, nnz * sizeof(I), cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(cols_h, cols, nnz * sizeof(I), cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(vals_h, vals, nnz * sizeof(T), cudaMemcpyDeviceToHost) != cudaSuccess) rc = RBX_ERR_CUDA;
    }
    cudaFree(rows); cudaFree(cols); cudaFree(vals);
    return rc;
}

// ---- dense-operator sketches ------------------------------------------------------------------------
// element strides (row, col) of a matrix stored in `layout` with leading dimension ld
static inline void strides_of(char layout, int64_t ld, int64_t* s_row, int64_t* s_col) {
    if (layout == 'C') { *s_row = 1; *s_col = ld; } else { *s_row = ld; *s_col = 1; }
}

template <typename T>
static int dispatch_dense_gemm(const GemmOperands<T>& g, cudaStream_t stream);
template <> int dispatch_dense_gemm<float>(const GemmOperands<float>& g, cudaStream_t stream) {
    int rc = launch_sketch_dense_gemv<float>(g, stream);
    if (rc != RBX_ERR_UNSUPPORTED) return rc;
    rc = launch_sketch_dense_tc_f32(g, stream);
    return rc == RBX_ERR_UNSUPPORTED ? launch_sketch_dense_simt<float>(g, stream) : rc;
}
template <> int dispatch_dense_gemm<double>(const GemmOperands<double>& g, cudaStream_t stream) {
    int rc = launch_sketch_dense_gemv<double>(g, stream);
    if (rc != RBX_ERR_UNSUPPORTED) return rc;
    rc = launch_sketch_dense_tc_f64(g, stream);
    return rc == RBX_ERR_UNSUPPORTED ? launch_sketch_dense_simt<double>(g, stream) : rc;
}

// Shared shape bookkeeping for both sides. The right sketch B = op(A) op(S) is run as B^T = op(S)^T op(A)^T.
struct SketchShape {
    int64_t M, N, K;            // output is M x N in "sketch-major" orientation: M = d always
    int64_t rows_sub, cols_sub; // dims of submat(S) before op
    bool s_trans;               // Sop[i, l] = S[ro + l, co + i] (true) or S[ro + i, co + l] (false)
    int64_t sc_i, sc_j, sx_l, sx_j;
};
static int sketch_shape(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2,
                        int64_t lda, int64_t ldb, bool check_lda, SketchShape* s) {
    RBX_REQUIRE(side == 'L' || side == 'R');
    RBX_REQUIRE(is_layout(layout) && is_op(opS) && is_op(opA));
    RBX_REQUIRE(p0 >= 0 && p1 >= 0 && p2 >= 0);
    int64_t a_r, a_c, b_r, b_c;
    strides_of(layout, lda, &a_r, &a_c);
    strides_of(layout, ldb, &b_r, &b_c);
    if (side == 'L') {
        const int64_t d = p0, n = p1, m = p2;
        s->M = d; s->N = n; s->K = m;
        s->rows_sub = (opS == 'N') ? d : m; s->cols_sub = (opS == 'N') ? m : d;
        s->s_trans = (opS == 'T');
        const int64_t rows_A = (opA == 'N') ? m : n, cols_A = (opA == 'N') ? n : m;
        if (layout == 'C') { if (check_lda) RBX_REQUIRE(lda >= rows_A); RBX_REQUIRE(ldb >= d); }
        else               { if (check_lda) RBX_REQUIRE(lda >= cols_A); RBX_REQUIRE(ldb >= n); }
        s->sc_i = b_r; s->sc_j = b_c;                            // C[i, j] = B[i, j]
        if (opA == 'N') { s->sx_l = a_r; s->sx_j = a_c; }        // X[l, j] = A[l, j]
        else            { s->sx_l = a_c; s->sx_j = a_r; }        // X[l, j] = A[j, l]
    } else {
        const int64_t m = p0, d = p1, n = p2;
        s->M = d; s->N = m; s->K = n;
        s->rows_sub = (opS == 'N') ? n : d; s->cols_sub = (opS == 'N') ? d : n;
        s->s_trans = (opS == 'N');                               // Sop[i, l] = op(S)[l, i]
        const int64_t rows_A = (opA == 'N') ? m : n, cols_A = (opA == 'N') ? n : m;
        if (layout == 'C') { if (check_lda) RBX_REQUIRE(lda >= rows_A); RBX_REQUIRE(ldb >= m); }
        else               { if (check_lda) RBX_REQUIRE(lda >= cols_A); RBX_REQUIRE(ldb >= d); }
        s->sc_i = b_c; s->sc_j = b_r;                            // C[i, j] = B[j, i]
        if (opA == 'N') { s->sx_l = a_c; s->sx_j = a_r; }        // X[l, j] = A[j, l]
        else            { s->sx_l = a_r; s->sx_j = a_c; }        // X[l, j] = A[l, j]
    }
    return RBX_OK;
}

template <typename T>
static int sketch_dense(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                        int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t* ctr, const uint32_t* key,
                        const T* S_buff, char S_layout, int64_t ro_s, int64_t co_s,
                        const T* A, int64_t lda, T beta, T* B, int64_t ldb, void* stream) {
    SketchShape s;
    RBX_TRY(sketch_shape(side, layout, opS, opA, p0, p1, p2, lda, ldb, true, &s));
    RBX_REQUIRE(ro_s >= 0 && co_s >= 0);
    RBX_REQUIRE(D_rows >= s.rows_sub + ro_s && D_cols >= s.cols_sub + co_s);   // skge.hh:197-198 / dense_skops.hh:488-489
    GemmOperands<T> g{};
    g.C = B; g.sc_i = s.sc_i; g.sc_j = s.sc_j;
    g.X = A; g.sx_l = s.sx_l; g.sx_j = s.sx_j;
    g.M = s.M; g.N = s.N; g.K = s.K;
    g.alpha = alpha; g.beta = beta;
    if (S_buff) {
        RBX_REQUIRE(is_layout(S_layout));
        int64_t s_r, s_c;
        strides_of(S_layout, S_layout == 'C' ? D_rows : D_cols, &s_r, &s_c);   // offset_and_ldim, base.hh:104-114
        g.Sx = S_buff + ro_s * s_r + co_s * s_c;
        g.ss_i = s.s_trans ? s_c : s_r;
        g.ss_l = s.s_trans ? s_r : s_c;
    } else {
        RBX_TRY(make_dense_window(D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s, &g.w));
        g.l_on_cols = ((!s.s_trans) == (g.w.parent_is_S != 0)) ? 1 : 0;
        g.Sx = nullptr;
    }
    RBX_NEED_DEVICE();
    return dispatch_dense_gemm<T>(g, (cudaStream_t) stream);
}

// sketch_vector without any allocation (rbx_sketch_vector_noalloc): the (RowMajor, n = 1) left sketch that
// skve.hh:152-176 forwards to, with the partial sums of the GEMV kernels in a static device array
template <typename T>
static int sketch_vector_noalloc(char opS, int64_t d, int64_t m, T alpha, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                                 const uint32_t* ctr, const uint32_t* key, int64_t ro_s, int64_t co_s,
                                 const T* x, int64_t incx, T beta, T* y, int64_t incy, void* stream) {
    RBX_REQUIRE(is_op(opS) && d >= 0 && m >= 0 && incx >= 1 && incy >= 1);
    const int64_t p0 = (opS == 'T') ? m : d, p2 = (opS == 'T') ? d : m;       // output / contraction length
    SketchShape s;
    RBX_TRY(sketch_shape('L', 'R', opS, 'N', p0, 1, p2, incx, incy, true, &s));
    RBX_REQUIRE(ro_s >= 0 && co_s >= 0);
    RBX_REQUIRE(D_rows >= s.rows_sub + ro_s && D_cols >= s.cols_sub + co_s);
    GemmOperands<T> g{};
    g.C = y; g.sc_i = s.sc_i; g.sc_j = s.sc_j;
    g.X = x; g.sx_l = s.sx_l; g.sx_j = s.sx_j;
    g.M = s.M; g.N = s.N; g.K = s.K;
    g.alpha = alpha; g.beta = beta;
    RBX_TRY(make_dense_window(D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s, &g.w));
    g.l_on_cols = ((!s.s_trans) == (g.w.parent_is_S != 0)) ? 1 : 0;
    g.Sx = nullptr;
    RBX_NEED_DEVICE();
    if (g.M == 0) return RBX_OK;
    if (g.K == 0) return launch_scale_view<T>(y, g.M, 1, g.sc_i, g.sc_j, beta, (cudaStream_t) stream);
    return launch_sketch_dense_gemv_static<T>(g, (cudaStream_t) stream);
}

// number of elements spanned by a (rows x cols) matrix in `layout` with leading dimension ld
static inline size_t span_elems(char layout, int64_t rows, int64_t cols, int64_t ld) {
    if (rows <= 0 || cols <= 0) return 0;
    return (layout == 'C') ? (size_t)((cols - 1) * ld + rows) : (size_t)((rows - 1) * ld + cols);
}

// How the staged copy of B returns to the host when B is a window of a larger array (ldb beyond the natural dimension):
// the kernels write only the rows x cols view, and the reference never touches the padding. Long unit-stride runs come back
// with a strided copy of the view only; short runs (sketch_vector with incy > 1 is the extreme: one element per run) upload
// the old contents first, so that the whole span can be copied back unchanged outside the view.
struct HostView {
    size_t span, inner, outer, ld;     // elements
    bool strided_back;                 // copy back the view with cudaMemcpy2D (otherwise the whole span)
    bool need_upload;                  // the padding of the device copy must hold the host's contents
};
template <typename T>
static inline HostView host_view(char layout, int64_t rows, int64_t cols, int64_t ld) {
    HostView v{};
    v.span = span_elems(layout, rows, cols, ld);
    v.inner = (size_t)(layout == 'C' ? rows : cols); v.outer = (size_t)(layout == 'C' ? cols : rows); v.ld = (size_t) ld;
    const bool padded = v.span != 0 && v.outer > 1 && v.ld > v.inner;
    v.strided_back = padded && v.inner * sizeof(T) >= 256;
    v.need_upload = padded && !v.strided_back;
    return v;
}
template <typename T>
static inline cudaError_t copy_back_view(T* host, const T* dev, const HostView& v, cudaStream_t st) {
    if (v.span == 0) return cudaSuccess;
    if (v.strided_back)
        return cudaMemcpy2DAsync(host, v.ld * sizeof(T), dev, v.ld * sizeof(T), v.inner * sizeof(T), v.outer, cudaMemcpyDeviceToHost, st);
    return cudaMemcpyAsync(host, dev, v.span * sizeof(T), cudaMemcpyDeviceToHost, st);
}

// Stage host A / B through device memory around `body(A_dev, B_dev, stream)`.
template <typename T, typename F>
static int with_staged(char layout, int64_t rows_A, int64_t cols_A, const T* A_host, int64_t lda,
                       int64_t rows_B, int64_t cols_B, T* B_host, int64_t ldb, bool read_B, F body) {
    { DryRun dry; RBX_TRY(body((const T*) nullptr, (T*) nullptr, nullptr)); }
    RBX_NEED_DEVICE();
    const size_t nA = span_elems(layout, rows_A, cols_A, lda);
    const HostView vb = host_view<T>(layout, rows_B, cols_B, ldb);
    const size_t nB = vb.span;
    read_B = read_B || vb.need_upload;
    T *A_dev = nullptr, *B_dev = nullptr;
    cudaStream_t st = nullptr;
    int rc = RBX_OK;
    // staging buffers come from the library's stream-ordered pool: a repeated call reuses the mapping of the previous one
    // instead of paying cudaMalloc / cudaFree (milliseconds for a few hundred MB) every time
    if ((nA && ws_alloc(&A_dev, nA * sizeof(T), st) != cudaSuccess) || (nB && ws_alloc(&B_dev, nB * sizeof(T), st) != cudaSuccess)) {
        cudaGetLastError(); rc = RBX_ERR_ALLOC;
    }
    if (rc == RBX_OK && nA && cudaMemcpyAsync(A_dev, A_host, nA * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK && nB && read_B && cudaMemcpyAsync(B_dev, B_host, nB * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK) rc = body(A_dev, B_dev, (void*) st);
    if (rc == RBX_OK && copy_back_view<T>(B_host, B_dev, vb, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (A_dev) cudaFreeAsync(A_dev, st);
    if (B_dev) cudaFreeAsync(B_dev, st);
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RBX_OK) rc = RBX_ERR_CUDA;
    return rc;
}

// ---- pageable host memory (what a caller's new T[] is) -----------------------------------------------------------------
// cudaMemcpy from pageable memory goes through the driver's small bounce buffer at a fraction of the link rate (measured at
// C2: 2.2 s per call against 0.32 s from pinned memory). The pipelined path below therefore stages pageable slabs itself:
// several host threads copy the slab into one of two pinned bounce buffers (kept for the life of the process; released by
// rbx_trim_workspace) while the previous slab crosses PCIe.
struct PinnedBounce {
    std::mutex mu;
    void* buf[2] = { nullptr, nullptr };
    size_t bytes = 0;
    bool get(size_t need) {
        if (need <= bytes && buf[0] && buf[1]) return true;
        release();
        for (int b = 0; b < 2; ++b)
            if (cudaHostAlloc(&buf[b], need, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); release(); return false; }
        bytes = need;
        return true;
    }
    void release() {
        for (int b = 0; b < 2; ++b) { if (buf[b]) cudaFreeHost(buf[b]); buf[b] = nullptr; }
        bytes = 0;
    }
};
static PinnedBounce& pinned_bounce() { static PinnedBounce p; return p; }

static bool is_pageable(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return at.type == cudaMemoryTypeUnregistered;
}
// dst (packed: `runs` runs of run_bytes) <- src (runs `src_pitch` bytes apart), split over host threads
static void parallel_gather_copy(uint8_t* dst, const uint8_t* src, size_t run_bytes, size_t runs, size_t src_pitch) {
    unsigned nt = std::thread::hardware_concurrency();
    if (nt > 16) nt = 16;
    if (nt < 1) nt = 1;
    const size_t total = run_bytes * runs;
    if (total < ((size_t) 8 << 20)) nt = 1;
    auto work = [=](unsigned t) {
        // thread t takes bytes [t, t + 1) * total / nt of the packed destination
        size_t b0 = total / nt * t, b1 = (t + 1 == nt) ? total : total / nt * (t + 1);
        while (b0 < b1) {
            const size_t run = b0 / run_bytes, off = b0 % run_bytes;
            size_t n = run_bytes - off;
            if (n > b1 - b0) n = b1 - b0;
            std::memcpy(dst + b0, src + run * src_pitch + off, n);
            b0 += n;
        }
    };
    if (nt == 1) { work(0); return; }
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
}

// Left sketch of a large host matrix, pipelined: op(A) = A (K x N) is cut into row slabs along the contraction axis; slab
// i+1 crosses PCIe while slab i is sketched with co_s advanced by the slab start and beta = 1 (the first slab carries the
// caller's beta). Device memory: two slabs instead of all of A. body(k0, klen, A_slab, lda_slab, beta_eff, B_dev, stream).
template <typename T, typename F>
static int staged_left_slabs(char layout, int64_t rows_B, int64_t N, int64_t K, const T* A_host, int64_t lda, T beta,
                             T* B_host, int64_t ldb, F body) {
    RBX_NEED_DEVICE();
    const HostView vb = host_view<T>(layout, rows_B, N, ldb);
    const size_t nB = vb.span;
    int64_t klen = (int64_t)((size_t) 512 << 20) / (N * (int64_t) sizeof(T));
    klen = (klen / 512) * 512;
    if (klen < 512) klen = 512;
    if (klen > K) klen = K;
    const int64_t nslab = (K + klen - 1) / klen;
    T* slab[2] = { nullptr, nullptr }; T* B_dev = nullptr;
    cudaStream_t sc = nullptr, sx = nullptr;
    cudaEvent_t copied[2] = { nullptr, nullptr }, consumed[2] = { nullptr, nullptr };
    int rc = RBX_OK;
    auto ok = [&](cudaError_t e) { if (e != cudaSuccess && rc == RBX_OK) { cudaGetLastError(); rc = RBX_ERR_CUDA; } return e == cudaSuccess; };
    // pageable A: stage through our own pinned bounce buffers (see PinnedBounce); pinned / registered A goes straight to the copy engine
    PinnedBounce& pb = pinned_bounce();
    std::unique_lock<std::mutex> pb_lock(pb.mu, std::defer_lock);
    bool bounce = is_pageable(A_host) && !getenv("RBX_NO_BOUNCE");
    if (bounce) { pb_lock.lock(); bounce = pb.get((size_t) klen * N * sizeof(T)); if (!bounce) pb_lock.unlock(); }
    if (ws_alloc(&slab[0], (size_t) klen * N * sizeof(T), nullptr) != cudaSuccess || ws_alloc(&slab[1], (size_t) klen * N * sizeof(T), nullptr) != cudaSuccess ||
        ws_alloc(&B_dev, nB * sizeof(T), nullptr) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_ALLOC; }
    if (rc == RBX_OK && cudaStreamSynchronize(nullptr) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }   // the buffers are used on other streams
    if (rc == RBX_OK) {
        ok(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking)); ok(cudaStreamCreateWithFlags(&sx, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) { ok(cudaEventCreateWithFlags(&copied[b], cudaEventDisableTiming)); ok(cudaEventCreateWithFlags(&consumed[b], cudaEventDisableTiming)); }
    }
    if (rc == RBX_OK && (beta != (T) 0 || vb.need_upload)) ok(cudaMemcpyAsync(B_dev, B_host, nB * sizeof(T), cudaMemcpyHostToDevice, sx));
    for (int64_t i = 0; i < nslab && rc == RBX_OK; ++i) {
        const int b = (int)(i & 1);
        const int64_t k0 = i * klen, kl = (k0 + klen < K) ? klen : K - k0;
        if (i >= 2) ok(cudaStreamWaitEvent(sc, consumed[b], 0));
        int64_t lda_slab;
        if (bounce) {
            // bounce buffer b was last read by the copy of slab i - 2: wait for that copy, pack the slab with host threads, then
            // one contiguous copy (the host packs slab i + 1 while slab i crosses PCIe and slab i - 1 is sketched)
            if (i >= 2) ok(cudaEventSynchronize(copied[b]));
            uint8_t* dst = static_cast<uint8_t*>(pb.buf[b]);
            if (layout == 'C') { lda_slab = kl; parallel_gather_copy(dst, reinterpret_cast<const uint8_t*>(A_host + k0), (size_t) kl * sizeof(T), (size_t) N, (size_t) lda * sizeof(T)); }
            else               { lda_slab = N;  parallel_gather_copy(dst, reinterpret_cast<const uint8_t*>(A_host + k0 * lda), (size_t) N * sizeof(T), (size_t) kl, (size_t) lda * sizeof(T)); }
            ok(cudaMemcpyAsync(slab[b], dst, (size_t) kl * N * sizeof(T), cudaMemcpyHostToDevice, sc));
        } else if (layout == 'C') { // K x N column-major: slab = kl contiguous elements of each of the N columns
            lda_slab = kl;
            ok(cudaMemcpy2DAsync(slab[b], (size_t) kl * sizeof(T), A_host + k0, (size_t) lda * sizeof(T), (size_t) kl * sizeof(T), (size_t) N,
                                 cudaMemcpyHostToDevice, sc));
        } else {                    // row-major: kl rows of N elements
            lda_slab = N;
            ok(cudaMemcpy2DAsync(slab[b], (size_t) N * sizeof(T), A_host + k0 * lda, (size_t) lda * sizeof(T), (size_t) N * sizeof(T), (size_t) kl,
                                 cudaMemcpyHostToDevice, sc));
        }
        ok(cudaEventRecord(copied[b], sc));
        ok(cudaStreamWaitEvent(sx, copied[b], 0));
        if (rc == RBX_OK) rc = body(k0, kl, (const T*) slab[b], lda_slab, i == 0 ? beta : (T) 1, B_dev, (void*) sx);
        ok(cudaEventRecord(consumed[b], sx));
    }
    if (rc == RBX_OK) ok(copy_back_view<T>(B_host, B_dev, vb, sx));
    if (sx) ok(cudaStreamSynchronize(sx));
    if (sc) ok(cudaStreamSynchronize(sc));
    for (int b = 0; b < 2; ++b) { if (copied[b]) cudaEventDestroy(copied[b]); if (consumed[b]) cudaEventDestroy(consumed[b]); }
    if (sc) cudaStreamDestroy(sc);
    if (sx) cudaStreamDestroy(sx);
    for (T* q : { slab[0], slab[1], B_dev }) if (q) cudaFreeAsync(q, nullptr);       // both work streams were synchronised above
    return rc;
}
// the pipelined path applies to plain left sketches of a host matrix of at least 1 GiB
template <typename T>
static inline bool use_slabs(char side, char opS, char opA, int64_t N, int64_t K) {
    return side == 'L' && opS == 'N' && opA == 'N' && (double) N * (double) K * sizeof(T) >= (double)((size_t) 1 << 30) && !getenv("RBX_NO_SLABS");
}

static inline void host_dims(char side, char opA, int64_t p0, int64_t p1, int64_t p2,
                             int64_t* rows_A, int64_t* cols_A, int64_t* rows_B, int64_t* cols_B) {
    if (side == 'L') { const int64_t d = p0, n = p1, m = p2; *rows_A = (opA == 'N') ? m : n; *cols_A = (opA == 'N') ? n : m; *rows_B = d; *cols_B = n; }
    else             { const int64_t m = p0, d = p1, n = p2; *rows_A = (opA == 'N') ? m : n; *cols_A = (opA == 'N') ? n : m; *rows_B = m; *cols_B = d; }
}

// ---- sparse-operator sketches -----------------------------------------------------------------------
template <typename T, typename I>
static int spmm_coo(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                    int64_t S_rows, int64_t S_cols, int64_t nnz, const I* rows, const I* cols, const T* vals,
                    int64_t ro_s, int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb, void* stream) {
    SketchShape s;
    RBX_TRY(sketch_shape(side, layout, opS, opA, p0, p1, p2, lda, ldb, true, &s));
    RBX_REQUIRE(ro_s >= 0 && co_s >= 0 && nnz >= 0);
    RBX_REQUIRE(S_rows >= s.rows_sub && S_cols >= s.cols_sub);               // spmm_dispatch.hh:94-96 (COO: >=)
    RBX_NEED_DEVICE();
    cudaStream_t st = (cudaStream_t) stream;
    // beta scaling (spmm_dispatch.hh:122-132) and the alpha == 0 early exit (:134-135) are part of the gather's single store
    return launch_coo_gather<T>(nnz, rows, cols, vals, s.s_trans, ro_s, co_s, s.M, s.N, s.K, alpha,
                                A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
}

// left_spmm / right_spmm with a CSR ('R') or CSC ('C') operand (spmm_dispatch.hh:48-200; the reference takes compressed operands
// whole, :97-104). When the operand is compressed along the output axis of the product -- CSR not transposed, CSC transposed --
// its arrays are used as they are: no regrouping, no copy. Otherwise the pointer array is expanded to one outer index per
// nonzero and the triplets take the regrouping path of spmm_coo.
template <typename T>
static int spmm_csx(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                    int64_t S_rows, int64_t S_cols, int64_t nnz, char fmt, const int64_t* ptr, const int64_t* idx, const T* vals,
                    const T* A, int64_t lda, T beta, T* B, int64_t ldb, void* stream) {
    SketchShape s;
    RBX_TRY(sketch_shape(side, layout, opS, opA, p0, p1, p2, lda, ldb, true, &s));
    RBX_REQUIRE((fmt == 'R' || fmt == 'C') && nnz >= 0);
    RBX_REQUIRE(S_rows == s.rows_sub && S_cols == s.cols_sub);               // spmm_dispatch.hh:97-104 (compressed: ==, no offsets)
    RBX_NEED_DEVICE();
    cudaStream_t st = (cudaStream_t) stream;
    const bool direct = (fmt == 'R') ? !s.s_trans : s.s_trans;
    if (direct)
        return launch_csx_gather<T>(ptr, idx, vals, nnz, s.M, s.N, s.K, alpha, A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
    if (nnz == 0 || alpha == (T) 0) return launch_scale_view<T>(B, s.M, s.N, s.sc_i, s.sc_j, beta, st);
    int64_t* outer = nullptr;
    if (ws_alloc(&outer, (size_t) nnz * 8, st) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    int rc = launch_expand_ptr(fmt == 'R' ? S_rows : S_cols, ptr, outer, st);
    if (rc == RBX_OK)
        rc = launch_coo_gather<T>(nnz, fmt == 'R' ? outer : idx, fmt == 'R' ? idx : outer, vals, s.s_trans, 0, 0, s.M, s.N, s.K, alpha,
                                  A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
    cudaFreeAsync(outer, st);
    return rc;
}

template <typename T>
static int sketch_sparseop(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                           int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t* ctr, const uint32_t* key,
                           int64_t ro_s, int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb, void* stream) {
    SketchShape s;
    RBX_TRY(sketch_shape(side, layout, opS, opA, p0, p1, p2, lda, ldb, true, &s));
    RBX_REQUIRE(major_axis == 'L' || major_axis == 'S');
    RBX_REQUIRE(S_rows > 0 && S_cols > 0 && vec_nnz > 0 && ro_s >= 0 && co_s >= 0);
    RBX_REQUIRE(S_rows >= s.rows_sub && S_cols >= s.cols_sub);
    const int64_t lng = S_rows > S_cols ? S_rows : S_cols, sht = S_rows < S_cols ? S_rows : S_cols;
    RBX_REQUIRE(vec_nnz <= (major_axis == 'S' ? sht : lng));
    RBX_NEED_DEVICE();
    cudaStream_t st = (cudaStream_t) stream;
    // streaming kernel for the SASO hot case; everything else regenerates the triplets and scatters
    if (!s.s_trans && major_axis == 'S' && S_rows <= S_cols) {
        // both data layouts are handled natively (rows of op(A) contiguous, or columns contiguous)
        const int rc = launch_saso_fast<T>(mk_ctr(ctr), mk_key(key), S_rows, S_cols, vec_nnz, ro_s, co_s, s.M, s.N, s.K, alpha,
                                           A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
        if (rc != RBX_ERR_UNSUPPORTED) return rc;
    }
    // A tall Short-axis operator used transposed (the usual right sketch B = A * S, S n x d) is the wide operator of the
    // transposed shape with the same state: fill_sparse draws one Fisher-Yates vector over the short axis per long-axis
    // index either way (sparse_skops.hh:392-410), only the roles of rows and columns swap. Square operators count as wide
    // there (:394), so their transpose is not of this form.
    if (s.s_trans && major_axis == 'S' && S_rows > S_cols) {
        const int rc = launch_saso_fast<T>(mk_ctr(ctr), mk_key(key), S_cols, S_rows, vec_nnz, co_s, ro_s, s.M, s.N, s.K, alpha,
                                           A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
        if (rc != RBX_ERR_UNSUPPORTED) return rc;
    }
    if (alpha == (T) 0) return launch_scale_view<T>(B, s.M, s.N, s.sc_i, s.sc_j, beta, st);
    // general path: regenerate the triplets (fill_sparse), then the deterministic output-stationary product; nothing is
    // written to B before every workspace has been allocated
    const size_t nnz = (size_t)(vec_nnz * (major_axis == 'S' ? lng : sht));
    int64_t *rows = nullptr, *cols = nullptr; T* vals = nullptr;
    if (ws_alloc(&rows, nnz * 8, st) != cudaSuccess || ws_alloc(&cols, nnz * 8, st) != cudaSuccess ||
        ws_alloc(&vals, nnz * sizeof(T), st) != cudaSuccess) {
        cudaGetLastError();
        if (rows) cudaFreeAsync(rows, st);
        if (cols) cudaFreeAsync(cols, st);
        return RBX_ERR_ALLOC;
    }
    int rc = fill_sparse<T, int64_t>(S_rows, S_cols, vec_nnz, major_axis, ctr, key, rows, cols, vals, nullptr, stream);
    if (rc == RBX_OK)
        rc = launch_coo_gather<T>((int64_t) nnz, rows, cols, vals, s.s_trans, ro_s, co_s, s.M, s.N, s.K, alpha,
                                  A, s.sx_l, s.sx_j, beta, B, s.sc_i, s.sc_j, st);
    cudaFreeAsync(rows, st); cudaFreeAsync(cols, st); cudaFreeAsync(vals, st);
    return rc;
}

// ---- dense operator x sparse data ---------------------------------------------------------------------
// Column blocks of the result as they finish (host-pointer calls copy block b back while block b + 1 is gathered). Only the
// one-pass schedule with B's unit stride along the sketch dimension uses it; `used` tells the caller whether it did.
struct ColBlockHook {
    int nblocks; bool used; void* ctx;
    int (*after)(void* ctx, int64_t j0, int64_t j1, cudaStream_t st);
};

template <typename T>
static int sketch_sparsedata(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                             int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t* ctr, const uint32_t* key,
                             int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                             const T* vals, const int64_t* idx0, const int64_t* idx1, int64_t ro_a, int64_t co_a,
                             T beta, T* B, int64_t ldb, void* stream, ColBlockHook* hook = nullptr) {
    SketchShape s;
    RBX_TRY(sketch_shape(side, layout, opS, opA, p0, p1, p2, 0, ldb, false, &s));
    RBX_REQUIRE(fmt == 'O' || fmt == 'R' || fmt == 'C');
    RBX_REQUIRE(ro_s >= 0 && co_s >= 0 && ro_a >= 0 && co_a >= 0 && nnz >= 0);
    RBX_REQUIRE(D_rows >= s.rows_sub + ro_s && D_cols >= s.cols_sub + co_s);     // sksp.hh:177-178
    // op(submat(A)) is K x N (left: m x n) or N x K (right: m x n with m = N, n = K)
    const int64_t a_m = (side == 'L') ? s.K : s.N, a_n = (side == 'L') ? s.N : s.K;
    const int64_t rows_sub_A = (opA == 'N') ? a_m : a_n, cols_sub_A = (opA == 'N') ? a_n : a_m;
    RBX_REQUIRE(A_rows >= rows_sub_A + ro_a && A_cols >= cols_sub_A + co_a);     // sksp.hh:175-176
    if (fmt != 'O') {                                                            // spmm_dispatch.hh:97-104
        RBX_REQUIRE(ro_a == 0 && co_a == 0);
        RBX_REQUIRE(A_rows == rows_sub_A && A_cols == cols_sub_A);
    }
    DenseOpWindow w;
    RBX_TRY(make_dense_window(D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s, &w));
    const bool l_on_cols = ((!s.s_trans) == (w.parent_is_S != 0));
    RBX_NEED_DEVICE();
    cudaStream_t st = (cudaStream_t) stream;
    if (alpha == (T) 0 || nnz == 0 || s.M == 0 || s.N == 0 || s.K == 0)
        return launch_scale_view<T>(B, s.M, s.N, s.sc_i, s.sc_j, beta, st);

    // entry (p, q) of A sits at (l, j) of op(A) for the left sketch when opA == 'N'; the right sketch is
    // transposed, so there (p, q) -> (j, l) when opA == 'N'.
    const bool a_trans = (side == 'L') ? (opA == 'T') : (opA == 'N');
    // column segments of op(A): CSC NoTrans and CSR Trans are grouped by j already (whole matrix, offsets 0);
    // everything else is regrouped on the device (rbx_spdata.cu)
    const int64_t *seg_ptr = nullptr, *seg_l = nullptr; const T* seg_v = nullptr;
    int64_t *own_ptr = nullptr, *own_l = nullptr, *expanded = nullptr; T* own_v = nullptr;
    int rc = RBX_OK;
    if (fmt == 'C' && !a_trans)      { seg_ptr = idx1; seg_l = idx0; seg_v = vals; }
    else if (fmt == 'R' && a_trans)  { seg_ptr = idx0; seg_l = idx1; seg_v = vals; }
    else {
        const int64_t *prow = idx0, *pcol = idx1;
        if (fmt != 'O') {
            if (ws_alloc(&expanded, (size_t) nnz * 8, st) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
            rc = (fmt == 'R') ? launch_expand_ptr(A_rows, idx0, expanded, st) : launch_expand_ptr(A_cols, idx1, expanded, st);
            if (fmt == 'R') { prow = expanded; pcol = idx1; } else { prow = idx0; pcol = expanded; }
        }
        if (rc == RBX_OK) rc = spd_group_by_column<T>(nnz, prow, pcol, vals, a_trans, ro_a, co_a, s.K, s.N, &own_ptr, &own_l, &own_v, st);
        if (expanded) cudaFreeAsync(expanded, st);
        if (rc != RBX_OK) return rc;
        seg_ptr = own_ptr; seg_l = own_l; seg_v = own_v;
    }
    const int64_t ma = major_axis_length(D_rows, D_cols, major_axis);
    const int64_t Mp = (s.M + 3) & ~(int64_t) 3;
    float* Sd = nullptr;
    // Two schedules. (1) One pass over the nonzeros, each gathering a full column of the float operator window
    // (spdata_gather_kernel): when the window (4 * Mp * K bytes) is far larger than L2 every gather is an HBM miss --
    // 198 GB of DRAM reads at C4, 36.4 ms. (2) RBX_SPD_DC = 8 / 16 / 32: the OUTPUT rows are blocked, dc at a time, with a
    // dc x K float slab (64 MB at C4 for dc = 16) generated per pass and kept in L2 (evict_last) while the nonzeros stream
    // through (evict_first). Measured at C4 (profiles/r02_c4_row_blocked_ncu.csv): DRAM reads fall to 81 GB, but the
    // gathers now run at the L2's bandwidth instead -- 407 GB of L2 traffic at 10.9 TB/s = 37.3 ms for the 32 passes, the
    // same time -- because with random sparsity nothing is reused above L2: nnz * 4 * M bytes must reach the SMs either
    // way. Schedule (1) therefore stays the default; (2) is kept, tested, for operators whose window exceeds device memory.
    int dc = 0;
    if (const char* ev = getenv("RBX_SPD_DC")) { const int v = atoi(ev); dc = (v == 8 || v == 16 || v == 32) ? v : 0; }
    if (dc == 0 && (double) Mp * (double) s.K * 4.0 > 0.5 * (double) ws_budget(~(size_t) 0 >> 1) * 4.0) dc = 16;   // window beyond half the free memory
    if (dc > 0) {
        float* slab[2] = { nullptr, nullptr };
        const size_t slab_bytes = 4 * (size_t) dc * (size_t) s.K;
        cudaStream_t aux = aux_stream();
        cudaEvent_t filled[2] = { nullptr, nullptr }, used[2] = { nullptr, nullptr };
        bool ok = ws_alloc(&slab[0], slab_bytes, st) == cudaSuccess && ws_alloc(&slab[1], slab_bytes, st) == cudaSuccess && aux != nullptr;
        for (int b = 0; b < 2 && ok; ++b)
            ok = cudaEventCreateWithFlags(&filled[b], cudaEventDisableTiming) == cudaSuccess && cudaEventCreateWithFlags(&used[b], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) { cudaGetLastError(); rc = RBX_ERR_ALLOC; }
        // the slab of pass p + 1 is generated on the auxiliary stream while pass p gathers
        auto fill = [&](int64_t i0, int b, cudaStream_t fs) -> int {
            const int64_t rows = (s.M - i0 < dc) ? s.M - i0 : dc;
            if (l_on_cols) return launch_fill_window<float>(family, ma, slab[b], rows, s.K, w.pr0 + i0, w.pc0, w.ctr, w.key, 1, dc, false, fs);
            return launch_fill_window<float>(family, ma, slab[b], s.K, rows, w.pr0, w.pc0 + i0, w.ctr, w.key, dc, 1, false, fs);
        };
        if (rc == RBX_OK && cudaEventRecord(used[0], st) != cudaSuccess) rc = RBX_ERR_CUDA;       // orders the first fill after the allocations
        if (rc == RBX_OK && cudaStreamWaitEvent(aux, used[0], 0) != cudaSuccess) rc = RBX_ERR_CUDA;
        if (rc == RBX_OK) rc = fill(0, 0, aux);
        if (rc == RBX_OK && cudaEventRecord(filled[0], aux) != cudaSuccess) rc = RBX_ERR_CUDA;
        int64_t pass = 0;
        for (int64_t i0 = 0; i0 < s.M && rc == RBX_OK; i0 += dc, ++pass) {
            const int b = (int)(pass & 1), nb = b ^ 1;
            const int64_t rows = (s.M - i0 < dc) ? s.M - i0 : dc;
            if (i0 + dc < s.M) {                       // next slab: after the gather that last read that buffer
                if (pass >= 1 && cudaStreamWaitEvent(aux, used[nb], 0) != cudaSuccess) rc = RBX_ERR_CUDA;
                if (rc == RBX_OK) rc = fill(i0 + dc, nb, aux);
                if (rc == RBX_OK && cudaEventRecord(filled[nb], aux) != cudaSuccess) rc = RBX_ERR_CUDA;
            }
            if (rc == RBX_OK && cudaStreamWaitEvent(st, filled[b], 0) != cudaSuccess) rc = RBX_ERR_CUDA;
            if (rc == RBX_OK) rc = launch_spdata_gather_narrow<T>(s.N, seg_ptr, seg_l, seg_v, s.K, (int) rows, dc, slab[b], family == 'U', alpha, beta,
                                                                 B + i0 * s.sc_i, s.sc_i, s.sc_j, st);
            if (rc == RBX_OK && cudaEventRecord(used[b], st) != cudaSuccess) rc = RBX_ERR_CUDA;
        }
        if (aux && cudaEventRecord(used[0], aux) == cudaSuccess) cudaStreamWaitEvent(st, used[0], 0);   // the caller's stream outlives the aux work
        for (int b = 0; b < 2; ++b) { if (filled[b]) cudaEventDestroy(filled[b]); if (used[b]) cudaEventDestroy(used[b]); if (slab[b]) cudaFreeAsync(slab[b], st); }
        if (rc != RBX_OK) cudaGetLastError();
    } else {
    // the operator window is generated in slabs of the contraction axis, as floats, Mp x Ks at a time (i contiguous)
    const int64_t budget = (int64_t) ws_budget((size_t) 4 << 30);       // at most 4 GiB, at most a quarter of the free memory
    int64_t Ks = budget / (4 * Mp);
    if (const char* ev = getenv("RBX_SPD_SLAB")) { const int64_t v = atoll(ev); if (v >= 1) Ks = v; }
    if (Ks < 1) Ks = 1;
    if (Ks > s.K) Ks = s.K;
    if (ws_alloc(&Sd, 4 * (size_t)(Mp * Ks), st) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_ALLOC; }
    for (int64_t l0 = 0; l0 < s.K && rc == RBX_OK; l0 += Ks) {
        const int64_t kk = (s.K - l0 < Ks) ? s.K - l0 : Ks;
        // unscaled floats: the uniform family's sqrt(3) is applied in T precision by the gather kernel
        if (l_on_cols) rc = launch_fill_window<float>(family, ma, Sd, s.M, kk, w.pr0, w.pc0 + l0, w.ctr, w.key, 1, Mp, false, st);
        else           rc = launch_fill_window<float>(family, ma, Sd, kk, s.M, w.pr0 + l0, w.pc0, w.ctr, w.key, Mp, 1, false, st);
        if (rc != RBX_OK) break;
        if (hook && hook->nblocks > 1 && Ks >= s.K && s.sc_i == 1) {      // one pass: the columns of B are final block by block
            hook->used = true;
            for (int b = 0; b < hook->nblocks && rc == RBX_OK; ++b) {
                const int64_t j0 = s.N * b / hook->nblocks, j1 = s.N * (b + 1) / hook->nblocks;
                rc = launch_spdata_gather<T>(j1 - j0, seg_ptr + j0, seg_l, seg_v, l0, kk, s.M, Mp, Sd, family == 'U', true, alpha, beta,
                                             B + j0 * s.sc_j, s.sc_i, s.sc_j, st);
                if (rc == RBX_OK) rc = hook->after(hook->ctx, j0, j1, st);
            }
            continue;
        }
        rc = launch_spdata_gather<T>(s.N, seg_ptr, seg_l, seg_v, l0, kk, s.M, Mp, Sd, family == 'U', l0 == 0, alpha, beta, B,
                                     s.sc_i, s.sc_j, st);
    }
    }
    if (Sd) cudaFreeAsync(Sd, st);
    if (own_ptr) cudaFreeAsync(own_ptr, st);
    if (own_l) cudaFreeAsync(own_l, st);
    if (own_v) cudaFreeAsync(own_v, st);
    return rc;
}

// sketch_sparse with every array in host memory (what a caller of the reference owns): the sparse matrix (vals and both index
// arrays) and B are staged through the library's stream-ordered pool around the device call; B comes back view-only when it
// is a window of a larger array. Argument errors are reported before the device is touched.
template <typename T>
static int host_sketch_sparsedata(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                                  int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t* ctr, const uint32_t* key,
                                  int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                                  const T* vals_h, const int64_t* idx0_h, const int64_t* idx1_h, int64_t ro_a, int64_t co_a,
                                  T beta, T* B_host, int64_t ldb) {
    { DryRun dry; RBX_TRY(sketch_sparsedata<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                                               fmt, A_rows, A_cols, nnz, (const T*) nullptr, (const int64_t*) nullptr, (const int64_t*) nullptr,
                                               ro_a, co_a, beta, (T*) nullptr, ldb, nullptr)); }
    RBX_NEED_DEVICE();
    const HostView vb = host_view<T>(layout, p0, p1, ldb);              // B is p0 x p1 on both sides (d x n left, m x d right)
    const bool read_B = beta != (T) 0 || vb.need_upload;
    const size_t n0 = (size_t)(fmt == 'R' ? A_rows + 1 : nnz), n1 = (size_t)(fmt == 'C' ? A_cols + 1 : nnz);
    T *vals = nullptr, *B_dev = nullptr; int64_t *i0 = nullptr, *i1 = nullptr;
    cudaStream_t st = nullptr;
    int rc = RBX_OK;
    if ((nnz && ws_alloc(&vals, (size_t) nnz * sizeof(T), st) != cudaSuccess) || (n0 && ws_alloc(&i0, n0 * 8, st) != cudaSuccess) ||
        (n1 && ws_alloc(&i1, n1 * 8, st) != cudaSuccess) || (vb.span && ws_alloc(&B_dev, vb.span * sizeof(T), st) != cudaSuccess)) {
        cudaGetLastError(); rc = RBX_ERR_ALLOC;
    }
    if (rc == RBX_OK && nnz && cudaMemcpyAsync(vals, vals_h, (size_t) nnz * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK && n0 && cudaMemcpyAsync(i0, idx0_h, n0 * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK && n1 && cudaMemcpyAsync(i1, idx1_h, n1 * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK && vb.span && read_B && cudaMemcpyAsync(B_dev, B_host, vb.span * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    // A large result goes back in column blocks on the auxiliary stream while the following blocks are still being gathered
    // (C4: 4.1 GB of B, 72 ms of PCIe against 36 ms of gather). j is B's outer axis whenever the hook is used.
    struct Back { T* host; const T* dev; HostView v; cudaStream_t copy; cudaEvent_t ev[16]; int nev; bool fail; } back{ B_host, B_dev, vb, nullptr, {}, 0, false };
    ColBlockHook hook{ 1, false, &back, [](void* c, int64_t j0, int64_t j1, cudaStream_t s2) -> int {
        Back& k = *static_cast<Back*>(c);
        if (j1 <= j0) return RBX_OK;
        cudaEvent_t e = nullptr;
        if (k.nev >= 16 || cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); k.fail = true; return RBX_ERR_CUDA; }
        k.ev[k.nev++] = e;
        bool ok = cudaEventRecord(e, s2) == cudaSuccess && cudaStreamWaitEvent(k.copy, e, 0) == cudaSuccess;
        const size_t off = (size_t) j0 * k.v.ld;
        if (k.v.strided_back)
            ok = ok && cudaMemcpy2DAsync(k.host + off, k.v.ld * sizeof(T), k.dev + off, k.v.ld * sizeof(T), k.v.inner * sizeof(T), (size_t)(j1 - j0),
                                         cudaMemcpyDeviceToHost, k.copy) == cudaSuccess;
        else {
            const size_t end = ((size_t) j1 * k.v.ld < k.v.span) ? (size_t) j1 * k.v.ld : k.v.span;
            ok = ok && cudaMemcpyAsync(k.host + off, k.dev + off, (end - off) * sizeof(T), cudaMemcpyDeviceToHost, k.copy) == cudaSuccess;
        }
        if (!ok) { cudaGetLastError(); k.fail = true; return RBX_ERR_CUDA; }
        return RBX_OK;
    } };
    size_t back_min = (size_t) 256 << 20;
    if (const char* ev = getenv("RBX_SPD_BACK_MB")) back_min = (size_t) atoll(ev) << 20;      // tests: block-wise copy-back at small sizes
    if (rc == RBX_OK && vb.span * sizeof(T) >= back_min && ((side == 'L') ? p1 : p0) >= 8 && !getenv("RBX_NO_SLABS") && (back.copy = aux_stream()) != nullptr) hook.nblocks = 8;
    if (rc == RBX_OK) rc = sketch_sparsedata<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                                                fmt, A_rows, A_cols, nnz, vals, i0, i1, ro_a, co_a, beta, B_dev, ldb, (void*) st, &hook);
    if (hook.used) {                                     // the caller's stream outlives the copies (B_dev is freed on it)
        cudaEvent_t done = nullptr;
        if (cudaEventCreateWithFlags(&done, cudaEventDisableTiming) != cudaSuccess || cudaEventRecord(done, back.copy) != cudaSuccess ||
            cudaStreamWaitEvent(st, done, 0) != cudaSuccess) { cudaGetLastError(); cudaStreamSynchronize(back.copy); if (rc == RBX_OK) rc = RBX_ERR_CUDA; }
        if (done) cudaEventDestroy(done);
        for (int b = 0; b < back.nev; ++b) cudaEventDestroy(back.ev[b]);
        if (back.fail && rc == RBX_OK) rc = RBX_ERR_CUDA;
    } else if (rc == RBX_OK && copy_back_view<T>(B_host, B_dev, vb, st) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (vals) cudaFreeAsync(vals, st);
    if (i0) cudaFreeAsync(i0, st);
    if (i1) cudaFreeAsync(i1, st);
    if (B_dev) cudaFreeAsync(B_dev, st);
    if (cudaStreamSynchronize(st) != cudaSuccess && rc == RBX_OK) rc = RBX_ERR_CUDA;
    return rc;
}

// ---- sparse format conversion -----------------------------------------------------------------------
template <typename T>
static int sparse_regroup(char by, char fmt_in, int64_t n_rows, int64_t n_cols, int64_t nnz, const T* vals, const int64_t* idx0,
                          const int64_t* idx1, int64_t* ptr_out, int64_t* idx_out, T* vals_out, void* stream) {
    RBX_REQUIRE(by == 'C' || by == 'R');
    RBX_REQUIRE(fmt_in == 'O' || fmt_in == 'R' || fmt_in == 'C');
    RBX_REQUIRE(n_rows >= 0 && n_cols >= 0 && nnz >= 0);
    RBX_NEED_DEVICE();
    cudaStream_t st = (cudaStream_t) stream;
    const int64_t n_outer = (by == 'C') ? n_cols : n_rows, n_inner = (by == 'C') ? n_rows : n_cols;
    const int64_t *prow = idx0, *pcol = idx1;
    int64_t* expanded = nullptr;
    int rc = RBX_OK;
    if (fmt_in != 'O' && nnz > 0) {
        if (ws_alloc(&expanded, (size_t) nnz * 8, st) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
        rc = (fmt_in == 'R') ? launch_expand_ptr(n_rows, idx0, expanded, st) : launch_expand_ptr(n_cols, idx1, expanded, st);
        if (fmt_in == 'R') { prow = expanded; pcol = idx1; } else { prow = idx0; pcol = expanded; }
    }
    int64_t *ptr = nullptr, *lidx = nullptr; T* v = nullptr;
    // grouping by columns keeps (l, j) = (row, col); grouping by rows swaps the roles
    if (rc == RBX_OK) rc = spd_group_by_column<T>(nnz, prow, pcol, vals, by == 'R', 0, 0, n_inner, n_outer, &ptr, &lidx, &v, st);
    if (expanded) cudaFreeAsync(expanded, st);
    if (rc != RBX_OK) return rc;
    bool ok = cudaMemcpyAsync(ptr_out, ptr, 8 * (size_t)(n_outer + 1), cudaMemcpyDeviceToDevice, st) == cudaSuccess;
    if (nnz > 0) ok = ok && cudaMemcpyAsync(idx_out, lidx, 8 * (size_t) nnz, cudaMemcpyDeviceToDevice, st) == cudaSuccess &&
                      cudaMemcpyAsync(vals_out, v, sizeof(T) * (size_t) nnz, cudaMemcpyDeviceToDevice, st) == cudaSuccess;
    cudaFreeAsync(ptr, st); cudaFreeAsync(lidx, st); cudaFreeAsync(v, st);
    if (!ok) { cudaGetLastError(); return RBX_ERR_CUDA; }
    return RBX_OK;
}

}  // namespace rbx

// =====================================================================================================
// extern "C" surface
// =====================================================================================================
using namespace rbx;

extern "C" {

const char* rbx_strerror(int code) {
    switch (code) {
        case RBX_OK: return "ok";
        case RBX_ERR_ARG: return "argument check failed (the reference would throw RandBLAS::exceptions::Error)";
        case RBX_ERR_CUDA: return "CUDA runtime or kernel launch failure";
        case RBX_ERR_NO_DEVICE: return "no sm_100 CUDA device visible (there is no CPU fallback)";
        case RBX_ERR_ALLOC: return "device workspace allocation failed";
        case RBX_ERR_UNSUPPORTED: return "case not implemented on the device";
        case RBX_ERR_COMM: return "NCCL unavailable (libnccl.so.2 not loadable) or an NCCL call failed";
        case RBX_ERR_TOO_LARGE: return "problem exceeds the fixed scratch capacity of an allocation-free entry point";
        default: return "unknown rbx error code";
    }
}
const char* rbx_version(void) { return "rbx 0.1 (sm_100a)"; }
int rbx_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; } return n; }
int64_t rbx_launch_count(void) { return g_launches.load(); }

void rbx_ctr_add(uint32_t ctr[4], uint64_t n) { put_ctr(ctr_add(mk_ctr(ctr), n), ctr); }
void rbx_dense_next_state(int64_t n_rows, int64_t n_cols, char major_axis, const uint32_t ctr[4], uint32_t next_ctr[4]) {
    const int64_t major = major_axis_length(n_rows, n_cols, major_axis);
    const int64_t minor = n_rows + (n_cols - major);
    put_ctr(ctr_add(mk_ctr(ctr), (uint64_t)(((major + 3) / 4) * minor)), next_ctr);
}
void rbx_sparse_next_state(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], uint32_t next_ctr[4]) {
    const int64_t lng = n_rows > n_cols ? n_rows : n_cols, sht = n_rows < n_cols ? n_rows : n_cols;
    put_ctr(ctr_add(mk_ctr(ctr), (uint64_t)(vec_nnz * (major_axis == 'S' ? sht : lng))), next_ctr);
}
char rbx_dist_to_layout(int64_t n_rows, int64_t n_cols, char major_axis) { return natural_is_rowmajor(n_rows, n_cols, major_axis) ? 'R' : 'C'; }
int64_t rbx_major_axis_length(int64_t n_rows, int64_t n_cols, char major_axis) { return major_axis_length(n_rows, n_cols, major_axis); }
void rbx_shard_range(int64_t total, int world, int rank, int64_t align, int64_t* start, int64_t* len) {
    if (align < 1) align = 1;
    const int64_t units = (total + align - 1) / align;
    const int64_t base = units / world, rem = units % world;
    const int64_t u0 = rank * base + (rank < rem ? rank : rem), nu = base + (rank < rem ? 1 : 0);
    int64_t s = u0 * align, e = (u0 + nu) * align;
    if (s > total) s = total;
    if (e > total) e = total;
    *start = s; *len = e - s;
}

int rbx_philox4x32_dev(const uint32_t ctr[4], const uint32_t key[2], int64_t n, uint32_t* out_dev, void* stream) {
    RBX_NEED_DEVICE();
    return launch_philox_stream(mk_ctr(ctr), mk_key(key), n, out_dev, (cudaStream_t) stream);
}
int rbx_generate4_dev(char family, const uint32_t ctr[4], const uint32_t key[2], int64_t n, float* out_dev, void* stream) {
    RBX_REQUIRE(family == 'G' || family == 'U');
    RBX_NEED_DEVICE();
    return launch_generate4_stream(family, mk_ctr(ctr), mk_key(key), n, out_dev, (cudaStream_t) stream);
}
int rbx_boxmuller_raw_dev(int64_t n, const uint32_t* u0_dev, const uint32_t* u1_dev, float* s_dev, float* c_dev, void* stream) {
    RBX_NEED_DEVICE();
    return launch_boxmuller_raw(n, u0_dev, u1_dev, s_dev, c_dev, (cudaStream_t) stream);
}

#define RBX_DEF_DENSE(SUF, T)                                                                                                     \
int rbx_fill_dense_submat_##SUF(char family, int64_t n_cols, T* smat_dev, int64_t n_srows, int64_t n_scols, int64_t ptr,          \
                                const uint32_t ctr[4], const uint32_t key[2], int64_t lda, uint32_t next_ctr[4], void* stream) {  \
    return fill_dense_submat<T>(family, n_cols, smat_dev, n_srows, n_scols, ptr, ctr, key, lda, next_ctr, stream);                \
}                                                                                                                                 \
int rbx_fill_dense_##SUF(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis, int64_t n_rows,               \
                         int64_t n_cols, int64_t ro_s, int64_t co_s, T* buff_dev, const uint32_t ctr[4], const uint32_t key[2],   \
                         uint32_t next_ctr[4], void* stream) {                                                                    \
    return fill_dense<T>(layout, D_rows, D_cols, family, major_axis, n_rows, n_cols, ro_s, co_s, buff_dev, ctr, key, next_ctr, stream); \
}                                                                                                                                 \
int rbx_host_fill_dense_##SUF(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis, int64_t n_rows,          \
                              int64_t n_cols, int64_t ro_s, int64_t co_s, T* buff_host, const uint32_t ctr[4],                    \
                              const uint32_t key[2], uint32_t next_ctr[4]) {                                                      \
    return host_fill_dense<T>(layout, D_rows, D_cols, family, major_axis, n_rows, n_cols, ro_s, co_s, buff_host, ctr, key, next_ctr); \
}                                                                                                                                 \
int rbx_sketch_dense_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,               \
                           int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4],                   \
                           const uint32_t key[2], const T* S_buff_dev, char S_layout, int64_t ro_s, int64_t co_s,                 \
                           const T* A_dev, int64_t lda, T beta, T* B_dev, int64_t ldb, void* stream) {                            \
    return sketch_dense<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key,               \
                           S_buff_dev, S_layout, ro_s, co_s, A_dev, lda, beta, B_dev, ldb, stream);                               \
}                                                                                                                                 \
int rbx_host_sketch_dense_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,          \
                                int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4],              \
                                const uint32_t key[2], int64_t ro_s, int64_t co_s, const T* A_host, int64_t lda, T beta,          \
                                T* B_host, int64_t ldb) {                                                                         \
    int64_t rA, cA, rB, cB;                                                                                                       \
    host_dims(side, opA, p0, p1, p2, &rA, &cA, &rB, &cB);                                                                         \
    if (use_slabs<T>(side, opS, opA, p1, p2)) {      /* >= 1 GiB of data: slab i + 1 crosses PCIe while slab i is sketched */     \
        { DryRun dry; RBX_TRY(sketch_dense<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, \
                                              key, nullptr, 'C', ro_s, co_s, (const T*) nullptr, lda, beta, (T*) nullptr, ldb, nullptr)); } \
        return staged_left_slabs<T>(layout, p0, p1, p2, A_host, lda, beta, B_host, ldb,                                           \
            [&](int64_t k0, int64_t kl, const T* A_slab, int64_t lda_slab, T beta_eff, T* B_dev, void* st) {                      \
                return sketch_dense<T>(side, layout, opS, opA, p0, p1, kl, alpha, D_rows, D_cols, family, major_axis, ctr, key,   \
                                       nullptr, 'C', ro_s, co_s + k0, A_slab, lda_slab, beta_eff, B_dev, ldb, st); });            \
    }                                                                                                                             \
    return with_staged<T>(layout, rA, cA, A_host, lda, rB, cB, B_host, ldb, beta != (T) 0,                                        \
        [&](const T* A_dev, T* B_dev, void* st) {                                                                                 \
            return sketch_dense<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key,       \
                                   nullptr, 'C', ro_s, co_s, A_dev, lda, beta, B_dev, ldb, st); });                               \
}                                                                                                                                 \
int rbx_sketch_sparseop_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,            \
                              int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4],            \
                              const uint32_t key[2], int64_t ro_s, int64_t co_s, const T* A_dev, int64_t lda, T beta,             \
                              T* B_dev, int64_t ldb, void* stream) {                                                              \
    return sketch_sparseop<T>(side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, vec_nnz, major_axis, ctr, key,           \
                              ro_s, co_s, A_dev, lda, beta, B_dev, ldb, stream);                                                  \
}                                                                                                                                 \
int rbx_host_sketch_sparseop_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,       \
                                   int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4],       \
                                   const uint32_t key[2], int64_t ro_s, int64_t co_s, const T* A_host, int64_t lda, T beta,       \
                                   T* B_host, int64_t ldb) {                                                                      \
    int64_t rA, cA, rB, cB;                                                                                                       \
    host_dims(side, opA, p0, p1, p2, &rA, &cA, &rB, &cB);                                                                         \
    if (use_slabs<T>(side, opS, opA, p1, p2)) {                                                                                   \
        { DryRun dry; RBX_TRY(sketch_sparseop<T>(side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, vec_nnz, major_axis,  \
                                                 ctr, key, ro_s, co_s, (const T*) nullptr, lda, beta, (T*) nullptr, ldb, nullptr)); } \
        return staged_left_slabs<T>(layout, p0, p1, p2, A_host, lda, beta, B_host, ldb,                                           \
            [&](int64_t k0, int64_t kl, const T* A_slab, int64_t lda_slab, T beta_eff, T* B_dev, void* st) {                      \
                return sketch_sparseop<T>(side, layout, opS, opA, p0, p1, kl, alpha, S_rows, S_cols, vec_nnz, major_axis, ctr,    \
                                          key, ro_s, co_s + k0, A_slab, lda_slab, beta_eff, B_dev, ldb, st); });                   \
    }                                                                                                                             \
    return with_staged<T>(layout, rA, cA, A_host, lda, rB, cB, B_host, ldb, beta != (T) 0,                                        \
        [&](const T* A_dev, T* B_dev, void* st) {                                                                                 \
            return sketch_sparseop<T>(side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, vec_nnz, major_axis, ctr, key,   \
                                      ro_s, co_s, A_dev, lda, beta, B_dev, ldb, st); });                                          \
}                                                                                                                                 \
int rbx_spmm_coo_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha, int64_t S_rows,   \
                       int64_t S_cols, int64_t nnz, const int64_t* rows_dev, const int64_t* cols_dev, const T* vals_dev,          \
                       int64_t ro_s, int64_t co_s, const T* A_dev, int64_t lda, T beta, T* B_dev, int64_t ldb, void* stream) {    \
    return spmm_coo<T, int64_t>(side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, nnz, rows_dev, cols_dev, vals_dev,     \
                                ro_s, co_s, A_dev, lda, beta, B_dev, ldb, stream);                                                \
}                                                                                                                                 \
int rbx_spmm_csx_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha, int64_t S_rows,   \
                       int64_t S_cols, int64_t nnz, char fmt, const int64_t* ptr_dev, const int64_t* idx_dev, const T* vals_dev,  \
                       const T* A_dev, int64_t lda, T beta, T* B_dev, int64_t ldb, void* stream) {                                \
    return spmm_csx<T>(side, layout, opS, opA, p0, p1, p2, alpha, S_rows, S_cols, nnz, fmt, ptr_dev, idx_dev, vals_dev,           \
                       A_dev, lda, beta, B_dev, ldb, stream);                                                                     \
}                                                                                                                                 \
int rbx_sketch_sparsedata_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,          \
                                int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4],              \
                                const uint32_t key[2], int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols,      \
                                int64_t nnz, const T* vals_dev, const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t ro_a,   \
                                int64_t co_a, T beta, T* B_dev, int64_t ldb, void* stream) {                                      \
    return sketch_sparsedata<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key,          \
                                ro_s, co_s, fmt, A_rows, A_cols, nnz, vals_dev, idx0_dev, idx1_dev, ro_a, co_a, beta, B_dev, ldb, stream); \
}                                                                                                                                 \
int rbx_host_sketch_sparsedata_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, T alpha,     \
                                     int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4],         \
                                     const uint32_t key[2], int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, \
                                     int64_t nnz, const T* vals_host, const int64_t* idx0_host, const int64_t* idx1_host,         \
                                     int64_t ro_a, int64_t co_a, T beta, T* B_host, int64_t ldb) {                                \
    return host_sketch_sparsedata<T>(side, layout, opS, opA, p0, p1, p2, alpha, D_rows, D_cols, family, major_axis, ctr, key,     \
                                     ro_s, co_s, fmt, A_rows, A_cols, nnz, vals_host, idx0_host, idx1_host, ro_a, co_a, beta,     \
                                     B_host, ldb);                                                                                \
}
RBX_DEF_DENSE(f32, float)
RBX_DEF_DENSE(f64, double)

#define RBX_DEF_SPARSE(SUF, T, I)                                                                                                 \
int rbx_repeated_fisher_yates_##SUF(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz, int64_t dim_major,             \
                                    int64_t dim_minor, I* idxs_major_dev, I* idxs_minor_dev, T* vals_dev, void* stream) {         \
    RBX_REQUIRE(vec_nnz <= dim_major);                                                                                            \
    RBX_NEED_DEVICE();                                                                                                            \
    return launch_fisher_yates<T, I>(mk_ctr(ctr), mk_key(key), vec_nnz, dim_major, dim_minor, idxs_major_dev, idxs_minor_dev,     \
                                     vals_dev, (cudaStream_t) stream);                                                            \
}                                                                                                                                 \
int rbx_fill_sparse_##SUF(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4],                \
                          const uint32_t key[2], I* rows_dev, I* cols_dev, T* vals_dev, uint32_t next_ctr[4], void* stream) {     \
    return fill_sparse<T, I>(n_rows, n_cols, vec_nnz, major_axis, ctr, key, rows_dev, cols_dev, vals_dev, next_ctr, stream);      \
}
RBX_DEF_SPARSE(f32_i32, float, int32_t)
RBX_DEF_SPARSE(f32_i64, float, int64_t)
RBX_DEF_SPARSE(f64_i32, double, int32_t)
RBX_DEF_SPARSE(f64_i64, double, int64_t)

#define RBX_DEF_SYM(SUF, T)                                                                                                       \
int rbx_require_symmetric_##SUF(char layout, const T* A_dev, int64_t n, int64_t lda, T tol, void* stream) {                       \
    RBX_REQUIRE(is_layout(layout) && n >= 0 && lda >= n);                                                                         \
    if (tol < (T) 0 || n <= 1) return RBX_OK;                                   /* util.hh:166: a negative tolerance skips the check */ \
    RBX_NEED_DEVICE();                                                                                                            \
    cudaStream_t st = (cudaStream_t) stream;                                                                                      \
    int* flag = nullptr;                                                                                                          \
    if (ws_alloc(&flag, sizeof(int), st) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }                            \
    int rc = cudaMemsetAsync(flag, 0, sizeof(int), st) == cudaSuccess ? RBX_OK : RBX_ERR_CUDA;                                    \
    int64_t s_row, s_col; strides_of(layout, lda, &s_row, &s_col);                                                                \
    if (rc == RBX_OK) rc = launch_require_symmetric<T>(A_dev, n, s_row, s_col, tol, flag, st);                                    \
    int host_flag = 0;                                                                                                            \
    if (rc == RBX_OK && (cudaMemcpyAsync(&host_flag, flag, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess ||             \
                         cudaStreamSynchronize(st) != cudaSuccess)) rc = RBX_ERR_CUDA;                                            \
    cudaFreeAsync(flag, st);                                                                                                      \
    if (rc != RBX_OK) return rc;                                                                                                  \
    return host_flag ? RBX_ERR_ARG : RBX_OK;                                                                                      \
}
RBX_DEF_SYM(f32, float)
RBX_DEF_SYM(f64, double)

int rbx_expand_ptr(int64_t n_outer, const int64_t* ptr_dev, int64_t* outer_idx_dev, void* stream) {
    RBX_REQUIRE(n_outer >= 0);
    RBX_NEED_DEVICE();
    return launch_expand_ptr(n_outer, ptr_dev, outer_idx_dev, (cudaStream_t) stream);
}
int rbx_sparse_regroup_f32(char by, char fmt_in, int64_t n_rows, int64_t n_cols, int64_t nnz, const float* vals_dev,
                           const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t* ptr_out_dev, int64_t* idx_out_dev,
                           float* vals_out_dev, void* stream) {
    return sparse_regroup<float>(by, fmt_in, n_rows, n_cols, nnz, vals_dev, idx0_dev, idx1_dev, ptr_out_dev, idx_out_dev, vals_out_dev, stream);
}
int rbx_sparse_regroup_f64(char by, char fmt_in, int64_t n_rows, int64_t n_cols, int64_t nnz, const double* vals_dev,
                           const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t* ptr_out_dev, int64_t* idx_out_dev,
                           double* vals_out_dev, void* stream) {
    return sparse_regroup<double>(by, fmt_in, n_rows, n_cols, nnz, vals_dev, idx0_dev, idx1_dev, ptr_out_dev, idx_out_dev, vals_out_dev, stream);
}

static void sample_next_ctr(const uint32_t ctr[4], int64_t k, uint32_t next_ctr[4]) {
    if (next_ctr) put_ctr(k > 0 ? ctr_add(mk_ctr(ctr), (uint64_t)((k + 3) / 4)) : mk_ctr(ctr), next_ctr);
}
int rbx_sample_indices_iid_uniform(int64_t n, int64_t k, int64_t* samples_dev, const uint32_t ctr[4], const uint32_t key[2],
                                   uint32_t next_ctr[4], void* stream) {
    RBX_REQUIRE(n >= 0 && k >= 0);
    sample_next_ctr(ctr, k, next_ctr);
    RBX_NEED_DEVICE();
    return launch_sample_uniform(n, k, samples_dev, mk_ctr(ctr), mk_key(key), (cudaStream_t) stream);
}
int rbx_sample_indices_iid_f32(int64_t n, const float* cdf_dev, int64_t k, int64_t* samples_dev, const uint32_t ctr[4],
                               const uint32_t key[2], uint32_t next_ctr[4], void* stream) {
    RBX_REQUIRE(n >= 0 && k >= 0);
    sample_next_ctr(ctr, k, next_ctr);
    RBX_NEED_DEVICE();
    return launch_sample_cdf<float>(n, cdf_dev, k, samples_dev, mk_ctr(ctr), mk_key(key), (cudaStream_t) stream);
}
int rbx_sample_indices_iid_f64(int64_t n, const double* cdf_dev, int64_t k, int64_t* samples_dev, const uint32_t ctr[4],
                               const uint32_t key[2], uint32_t next_ctr[4], void* stream) {
    RBX_REQUIRE(n >= 0 && k >= 0);
    sample_next_ctr(ctr, k, next_ctr);
    RBX_NEED_DEVICE();
    return launch_sample_cdf<double>(n, cdf_dev, k, samples_dev, mk_ctr(ctr), mk_key(key), (cudaStream_t) stream);
}
int rbx_weights_to_cdf_f32(int64_t n, float* w_dev, float error_if_below, void* stream) {
    RBX_REQUIRE(n >= 0);
    RBX_NEED_DEVICE();
    return launch_weights_to_cdf<float>(n, w_dev, error_if_below, (cudaStream_t) stream);
}
int rbx_weights_to_cdf_f64(int64_t n, double* w_dev, double error_if_below, void* stream) {
    RBX_REQUIRE(n >= 0);
    RBX_NEED_DEVICE();
    return launch_weights_to_cdf<double>(n, w_dev, error_if_below, (cudaStream_t) stream);
}

int rbx_sketch_vector_noalloc(char dtype, char opS, int64_t d, int64_t m, double alpha, int64_t D_rows, int64_t D_cols, char family,
                              char major_axis, const uint32_t ctr[4], const uint32_t key[2], int64_t ro_s, int64_t co_s,
                              const void* x_dev, int64_t incx, double beta, void* y_dev, int64_t incy, void* stream) {
    if (dtype == 'f')
        return sketch_vector_noalloc<float>(opS, d, m, (float) alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                                            (const float*) x_dev, incx, (float) beta, (float*) y_dev, incy, stream);
    if (dtype == 'd')
        return sketch_vector_noalloc<double>(opS, d, m, alpha, D_rows, D_cols, family, major_axis, ctr, key, ro_s, co_s,
                                             (const double*) x_dev, incx, beta, (double*) y_dev, incy, stream);
    return RBX_ERR_ARG;
}

int rbx_malloc(void** dev_ptr, uint64_t bytes) {
    if (!dev_ptr) return RBX_ERR_ARG;
    *dev_ptr = nullptr;
    RBX_NEED_DEVICE();
    if (bytes == 0) return RBX_OK;
    if (cudaMalloc(dev_ptr, (size_t) bytes) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    return RBX_OK;
}
int rbx_free(void* dev_ptr) { return (dev_ptr == nullptr || cudaFree(dev_ptr) == cudaSuccess) ? RBX_OK : RBX_ERR_CUDA; }
int rbx_memcpy_h2d(void* dst_dev, const void* src_host, uint64_t bytes, void* stream) {
    if (bytes == 0) return RBX_OK;
    return cudaMemcpyAsync(dst_dev, src_host, (size_t) bytes, cudaMemcpyHostToDevice, (cudaStream_t) stream) == cudaSuccess ? RBX_OK : RBX_ERR_CUDA;
}
int rbx_memcpy_d2h(void* dst_host, const void* src_dev, uint64_t bytes, void* stream) {
    if (bytes == 0) return RBX_OK;
    return cudaMemcpyAsync(dst_host, src_dev, (size_t) bytes, cudaMemcpyDeviceToHost, (cudaStream_t) stream) == cudaSuccess ? RBX_OK : RBX_ERR_CUDA;
}
int rbx_stream_synchronize(void* stream) { return cudaStreamSynchronize((cudaStream_t) stream) == cudaSuccess ? RBX_OK : RBX_ERR_CUDA; }
int rbx_trim_workspace(void) {
    int dev = 0;
    if (cudaDeviceSynchronize() != cudaSuccess || cudaGetDevice(&dev) != cudaSuccess) return RBX_ERR_CUDA;
    { PinnedBounce& pb = pinned_bounce(); std::lock_guard<std::mutex> g(pb.mu); pb.release(); }
    cudaMemPool_t pool = ws_pool(dev);
    if (!pool) return RBX_OK;
    return cudaMemPoolTrimTo(pool, 0) == cudaSuccess ? RBX_OK : RBX_ERR_CUDA;
}

int rbx_workspace_stats(uint64_t* used_now, uint64_t* used_high, uint64_t* reserved_now, int reset_high) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_NO_DEVICE; }
    cudaMemPool_t pool = ws_pool(dev);
    unsigned long long un = 0, uh = 0, rn = 0;
    if (pool) {
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &un) != cudaSuccess ||
            cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &uh) != cudaSuccess ||
            cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &rn) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_CUDA; }
        if (reset_high) {
            unsigned long long zero = 0;
            if (cudaMemPoolSetAttribute(pool, cudaMemPoolAttrUsedMemHigh, &zero) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_CUDA; }
        }
    }
    if (used_now) *used_now = un;
    if (used_high) *used_high = uh;
    if (reserved_now) *reserved_now = rn;
    return RBX_OK;
}

int rbx_host_fill_sparse_f64_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t* rows_host, int64_t* cols_host, double* vals_host, uint32_t next_ctr[4]) {
    return host_fill_sparse<double, int64_t>(n_rows, n_cols, vec_nnz, major_axis, ctr, key, rows_host, cols_host, vals_host, next_ctr);
}
int rbx_host_fill_sparse_f32_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t* rows_host, int64_t* cols_host, float* vals_host, uint32_t next_ctr[4]) {
    return host_fill_sparse<float, int64_t>(n_rows, n_cols, vec_nnz, major_axis, ctr, key, rows_host, cols_host, vals_host, next_ctr);
}

}  // extern "C"
