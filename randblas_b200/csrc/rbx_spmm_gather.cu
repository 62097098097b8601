// K4' general sparse-operator product, deterministic: C (M x N) = alpha * Sop (M x K) * X (K x N) + beta * C for an operator
// given as COO triplets (an explicit or caller-edited SparseSkOp, a LASO, a tall operator, a SASO outside the streaming
// kernel's envelope, left_spmm / right_spmm on a COO / CSR / CSC matrix).
//
// Replaces left_spmm -> coo::apply_coo_left_jki_p11 (spmm_dispatch.hh:48-160, coo_spmm_impl.hh:79-162), which sorts the
// triplets on the host and walks them once per column of X. Round 1 scattered with atomicAdd (run-to-run different sums);
// this is the output-stationary form:
//   * the entries are regrouped by output row i on the device (count, scan, scatter, rank sort: the kernels of
//     rbx_spdata.cu), each row cut into sub-segments by blocks of the contraction axis so that no sorted segment is long;
//     within a row the entries come out in increasing l (duplicates ordered by their value bits);
//   * rows of X contiguous: one warp owns 128 columns of one output row in registers and adds v * X[l, :] entry after entry
//     (coalesced 128-byte reads, entry metadata fetched 32 at a time and broadcast by shuffles);
//   * columns of X contiguous: one warp owns 4 columns of one output row; its lanes stride over the row's entries and the
//     32 partial sums are folded by a butterfly in a fixed order;
//   * alpha / beta are applied at the single store of each element (C is not read when beta == 0, util.hh:51-59).
// No atomics, fixed summation order => bitwise reproducible. Roofline: HBM (nnz rows of X re-read; L2 absorbs what fits).
#include "rbx_kernels.cuh"

namespace rbx {

template <typename T>
struct CooGatherArgs {
    const int64_t* ptr; const int64_t* lidx; const T* vals;     // sub-segments: row i owns [ptr[i * nlc], ptr[(i + 1) * nlc])
    int64_t nlc;
    int64_t M, N;
    const T* X; int64_t sx_l, sx_j;
    T alpha, beta;
    T* C; int64_t sc_i, sc_j;
};

// rows of X contiguous (sx_j == 1; also the fallback for general strides): grid = (ceil(N / 128), ceil(M / WPB)); one warp
// per (row i, 128-column tile), 4 accumulators per lane; the WPB warps of a block take consecutive rows of the same tile.
constexpr int CG_WPB = 8;
template <typename T>
__global__ void __launch_bounds__(32 * CG_WPB) coo_gather_rows_kernel(const CooGatherArgs<T> a) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i = (int64_t) blockIdx.y * CG_WPB + w;
    if (i >= a.M) return;
    const int64_t j0 = (int64_t) blockIdx.x * 128 + lane;
    const int64_t beg = a.ptr[i * a.nlc], end = a.ptr[(i + 1) * a.nlc];
    T acc[4] = { (T) 0, (T) 0, (T) 0, (T) 0 };
    bool live[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) live[c] = j0 + 32 * c < a.N;
    for (int64_t e0 = beg; e0 < end; e0 += 32) {
        int64_t my_l = 0; T my_v = (T) 0;
        if (e0 + lane < end) { my_l = a.lidx[e0 + lane]; my_v = a.vals[e0 + lane]; }
        const int cnt = (end - e0 < 32) ? (int)(end - e0) : 32;
        int u = 0;
        for (; u + 4 <= cnt; u += 4) {                           // four entries in flight
            T x[4][4], v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t l = __shfl_sync(0xffffffffu, my_l, u + q);
                v[q] = __shfl_sync(0xffffffffu, my_v, u + q);
                const T* row = a.X + l * a.sx_l + j0;
#pragma unroll
                for (int c = 0; c < 4; ++c) x[q][c] = live[c] ? __ldg(row + 32 * c) : (T) 0;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[c] = fma(v[q], x[q][c], acc[c]);
        }
        for (; u < cnt; ++u) {
            const int64_t l = __shfl_sync(0xffffffffu, my_l, u);
            const T v = __shfl_sync(0xffffffffu, my_v, u);
            const T* row = a.X + l * a.sx_l + j0;
#pragma unroll
            for (int c = 0; c < 4; ++c) if (live[c]) acc[c] = fma(v, __ldg(row + 32 * c), acc[c]);
        }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c)
        if (live[c]) {
            T* p = a.C + i * a.sc_i + (j0 + 32 * c) * a.sc_j;
            const T r = a.alpha * acc[c];
            *p = (a.beta == (T) 0) ? r : r + a.beta * (*p);
        }
}

// columns of X contiguous (sx_l == 1): grid = (ceil(N / 4), ceil(M / WPB)); a warp owns (row i, 4 columns); lanes stride over
// the entries (consecutive lanes read increasing l of the same column: the sectors of a sorted segment are shared)
template <typename T>
__global__ void __launch_bounds__(32 * CG_WPB) coo_gather_cols_kernel(const CooGatherArgs<T> a) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t i = (int64_t) blockIdx.y * CG_WPB + w;
    if (i >= a.M) return;
    const int64_t j0 = (int64_t) blockIdx.x * 4;
    const int64_t beg = a.ptr[i * a.nlc], end = a.ptr[(i + 1) * a.nlc];
    T acc[4] = { (T) 0, (T) 0, (T) 0, (T) 0 };
    const int nj = (a.N - j0 < 4) ? (int)(a.N - j0) : 4;
    for (int64_t e = beg + lane; e < end; e += 32) {
        const int64_t l = a.lidx[e];
        const T v = a.vals[e];
        const T* col = a.X + l * a.sx_l + j0 * a.sx_j;
#pragma unroll
        for (int c = 0; c < 4; ++c) if (c < nj) acc[c] = fma(v, __ldg(col + c * a.sx_j), acc[c]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], o);     // fixed butterfly: the same sum in every lane
    if (lane < nj) {
        T s = acc[0];
        if (lane == 1) s = acc[1];
        if (lane == 2) s = acc[2];
        if (lane == 3) s = acc[3];
        T* p = a.C + i * a.sc_i + (j0 + lane) * a.sc_j;
        const T r = a.alpha * s;
        *p = (a.beta == (T) 0) ? r : r + a.beta * (*p);
    }
}

// dst[l * pitch + jj] = X[l + (j0 + jj) * sx_j] for K x nc: a block of columns of column-contiguous data re-laid with contiguous rows
template <typename T>
__global__ void __launch_bounds__(256) cg_rows_copy_kernel(const T* __restrict__ X, int64_t K, int64_t nc, int64_t sx_j, T* __restrict__ dst, int64_t pitch) {
    __shared__ T tile[64][65];
    const int64_t l0 = (int64_t) blockIdx.x * 64, j0 = (int64_t) blockIdx.y * 64;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    for (int jj = ty; jj < 64; jj += 4) {
        const int64_t l = l0 + tx, j = j0 + jj;
        tile[jj][tx] = (l < K && j < nc) ? X[l + j * sx_j] : (T) 0;
    }
    __syncthreads();
    for (int ll = ty; ll < 64; ll += 4) {
        const int64_t l = l0 + ll, j = j0 + tx;
        if (l < K && j < nc) dst[l * pitch + j] = tile[tx][ll];
    }
}

// the gather kernels over ready-made segments: row i of the output owns entries [ptr[i * nlc], ptr[(i + 1) * nlc])
template <typename T>
static int run_gather_segments(const int64_t* ptr, const int64_t* lidx, const T* v, int64_t nlc, int64_t M, int64_t N, T alpha,
                               const T* X, int64_t sx_l, int64_t sx_j, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream,
                               int64_t K = 0, int64_t nnz = 0) {
    int rc = RBX_OK;
    // Column-contiguous data with many columns: a gather of row l touches one 8-byte element per 32-byte sector (a quarter of
    // the bytes moved are used) and the 4-column kernel below re-reads the entry list N / 4 times. When the product moves far
    // more bytes than the data holds (nnz >= K / 2: every sketching operator), blocks of columns are first re-laid with
    // contiguous rows into a workspace of at most 1 GB and gathered with full 128-byte rows (measured, d = 2048, m = 2^18,
    // n = 2048, double: LASO with 512 entries per row 41 -> 7.2 ms; SASO with vec_nnz = 40 230 -> 16.9 ms).
    if (sx_l == 1 && sx_j != 1 && N >= 128 && K > 0 && nnz >= K / 2 && K <= 65535ll * 64 && !getenv("RBX_CG_NO_RELAYOUT")) {
        int64_t budget = (int64_t) 1 << 30;
        if (const char* ev = getenv("RBX_CG_RELAYOUT_MB")) { const int64_t mb = atoll(ev); if (mb >= 1) budget = mb << 20; }
        int64_t Nc = budget / (K * (int64_t) sizeof(T)) / 128 * 128;
        if (Nc >= 128) {
            if (Nc > N) Nc = (N + 127) / 128 * 128;
            T* wsx = nullptr;
            if (ws_alloc(&wsx, sizeof(T) * (size_t) K * (size_t) Nc, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
            for (int64_t j0 = 0; j0 < N && rc == RBX_OK; j0 += Nc) {
                const int64_t nc = (N - j0 < Nc) ? N - j0 : Nc;
                cg_rows_copy_kernel<T><<<dim3((unsigned)((K + 63) / 64), (unsigned)((nc + 63) / 64)), 256, 0, stream>>>(X + j0 * sx_j, K, nc, sx_j, wsx, Nc);
                count_launch();
                rc = check_launch("cg_rows_copy_kernel");
                if (rc == RBX_OK) rc = run_gather_segments<T>(ptr, lidx, v, nlc, M, nc, alpha, wsx, Nc, 1, beta, C + j0 * sc_j, sc_i, sc_j, stream);
            }
            cudaFreeAsync(wsx, stream);
            return rc;
        }
    }
    CooGatherArgs<T> a;
    a.ptr = ptr; a.lidx = lidx; a.vals = v; a.nlc = nlc; a.M = M; a.N = N; a.X = X; a.sx_l = sx_l; a.sx_j = sx_j;
    a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j;
    const int64_t by = (M + CG_WPB - 1) / CG_WPB;
    // grid.y is limited to 65535: tile the rows in launches of that many blocks
    for (int64_t y0 = 0; y0 < by && rc == RBX_OK; y0 += 65535) {
        const int64_t ny = (by - y0 < 65535) ? by - y0 : 65535;
        CooGatherArgs<T> b = a;
        b.ptr = a.ptr + y0 * CG_WPB * nlc; b.M = (M - y0 * CG_WPB < ny * CG_WPB) ? M - y0 * CG_WPB : ny * CG_WPB;
        b.C = a.C + y0 * CG_WPB * sc_i;
        if (sx_j == 1 || sx_l != 1) {
            const int64_t bx = (N + 127) / 128;
            if (bx > 0x7fffffffLL) { rc = RBX_ERR_UNSUPPORTED; break; }
            coo_gather_rows_kernel<T><<<dim3((unsigned) bx, (unsigned) ny), 32 * CG_WPB, 0, stream>>>(b);
        } else {
            const int64_t bx = (N + 3) / 4;
            if (bx > 0x7fffffffLL) { rc = RBX_ERR_UNSUPPORTED; break; }
            coo_gather_cols_kernel<T><<<dim3((unsigned) bx, (unsigned) ny), 32 * CG_WPB, 0, stream>>>(b);
        }
        count_launch();
        rc = check_launch("coo_gather_kernel");
    }
    return rc;
}

template <typename T>
int launch_coo_gather(int64_t nnz, const int64_t* rows, const int64_t* cols, const T* vals, bool trans, int64_t ro, int64_t co,
                      int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                      T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return RBX_OK;
    if (alpha == (T) 0 || nnz <= 0 || K <= 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);
    // sub-segments: aim at ~64 entries each (the rank sort works out of shared memory up to 256), at most 2^22 segments
    int lc_shift = 62; int64_t nlc = 1;
    const double per_row = (double) nnz / (double) M;
    if (per_row > 128.0) {
        int64_t want = (int64_t)(per_row / 64.0);
        const int64_t cap = ((int64_t) 1 << 22) / (M > 0 ? M : 1);
        if (want > cap) want = cap;
        if (want < 1) want = 1;
        int sh = 0;
        while (((K + ((int64_t) 1 << sh) - 1) >> sh) > want) ++sh;
        lc_shift = sh; nlc = (K + ((int64_t) 1 << sh) - 1) >> sh;
        if (nlc < 1) nlc = 1;
    }
    // group by output row: the regroup kernels map (p, q) -> (l, j) = trans' ? (q - co, p - ro) : (p - ro, q - co) and
    // group by j; here j must be the output row i, so trans' = !trans (see rbx_kernels.cuh)
    int64_t *ptr = nullptr, *lidx = nullptr; T* v = nullptr;
    int rc = spd_group_by_column<T>(nnz, rows, cols, vals, !trans, ro, co, K, M, &ptr, &lidx, &v, stream, lc_shift, nlc);
    if (rc != RBX_OK) return rc;
    rc = run_gather_segments<T>(ptr, lidx, v, nlc, M, N, alpha, X, sx_l, sx_j, beta, C, sc_i, sc_j, stream, K, nnz);
    cudaFreeAsync(ptr, stream); cudaFreeAsync(lidx, stream); cudaFreeAsync(v, stream);
    return rc;
}
template int launch_coo_gather<float>(int64_t, const int64_t*, const int64_t*, const float*, bool, int64_t, int64_t, int64_t, int64_t, int64_t,
                                      float, const float*, int64_t, int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_coo_gather<double>(int64_t, const int64_t*, const int64_t*, const double*, bool, int64_t, int64_t, int64_t, int64_t, int64_t,
                                       double, const double*, int64_t, int64_t, double, double*, int64_t, int64_t, cudaStream_t);

// The same product for an operand that is already compressed along the output axis (CSR used as is, CSC used transposed):
// its pointer / index / value arrays ARE the segments, so nothing is regrouped or copied. Entries of a row are summed in
// storage order (fixed, so the result is still bitwise reproducible).
template <typename T>
int launch_csx_gather(const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t nnz, int64_t M, int64_t N, int64_t K, T alpha,
                      const T* X, int64_t sx_l, int64_t sx_j, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return RBX_OK;
    if (alpha == (T) 0 || nnz <= 0 || K <= 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);
    return run_gather_segments<T>(ptr, lidx, vals, 1, M, N, alpha, X, sx_l, sx_j, beta, C, sc_i, sc_j, stream, K, nnz);
}
template int launch_csx_gather<float>(const int64_t*, const int64_t*, const float*, int64_t, int64_t, int64_t, int64_t, float, const float*, int64_t,
                                      int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_csx_gather<double>(const int64_t*, const int64_t*, const double*, int64_t, int64_t, int64_t, int64_t, double, const double*, int64_t,
                                       int64_t, double, double*, int64_t, int64_t, cudaStream_t);

}  // namespace rbx
