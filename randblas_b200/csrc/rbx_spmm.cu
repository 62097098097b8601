// K4/K5 (general paths): sparse-times-dense products behind lskges/rskges and lsksp3/rsksp3.
//
// Replaces left_spmm / right_spmm (spmm_dispatch.hh:48-200) and the OpenMP kernels they select
// (coo_spmm_impl.hh:79-162, csc_spmm_impl.hh:43-213, csr_spmm_impl.hh:45-156). The reference sorts the
// COO triplets into CSC order on the host and walks one column of the dense operand per OpenMP task;
// these kernels take the triplets in any order, apply submatrix filtering and transposition on the fly
// (set_filtered_coo, coo_spmm_impl.hh:45-75) and accumulate with red.global.add. They are the
// shape-agnostic fallbacks; rbx_saso.cu holds the conflict-free streaming kernel for the SASO hot case.
// Both are HBM/L2-bound byte-moving kernels: no tensor cores involved.
#include "rbx_kernels.cuh"

namespace rbx {

template <typename T>
__global__ void __launch_bounds__(256) scale_view_kernel(T* __restrict__ C, int64_t M, int64_t N, int64_t sc_i, int64_t sc_j, T beta,
                                                        bool i_fastest) {
    const int64_t total = M * N;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        int64_t i, j;
        if (i_fastest) { j = e / M; i = e - j * M; } else { i = e / N; j = e - i * N; }
        T* c = C + i * sc_i + j * sc_j;
        *c = (beta == (T) 0) ? (T) 0 : beta * (*c);
    }
}

template <typename T>
int launch_scale_view(T* C, int64_t M, int64_t N, int64_t sc_i, int64_t sc_j, T beta, cudaStream_t stream) {
    if (M <= 0 || N <= 0 || beta == (T) 1) return RBX_OK;
    int64_t blocks = (M * N + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    scale_view_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(C, M, N, sc_i, sc_j, beta, sc_i <= sc_j);
    count_launch();
    return check_launch("scale_view_kernel");
}
template int launch_scale_view<float>(float*, int64_t, int64_t, int64_t, int64_t, float, cudaStream_t);
template int launch_scale_view<double>(double*, int64_t, int64_t, int64_t, int64_t, double, cudaStream_t);

// One warp per entry, lanes along the columns of X / C (use when rows of X are contiguous).
template <typename T, typename I>
__global__ void __launch_bounds__(256) coo_scatter_rows_kernel(int64_t nnz, const I* __restrict__ rows, const I* __restrict__ cols,
                                                              const T* __restrict__ vals, bool trans, int64_t ro, int64_t co,
                                                              int64_t M, int64_t N, int64_t K, T alpha, const T* __restrict__ X,
                                                              int64_t sx_l, int64_t sx_j, T* __restrict__ C, int64_t sc_i, int64_t sc_j) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t e = warp; e < nnz; e += nwarps) {
        const int64_t r = (int64_t) rows[e] - ro, c = (int64_t) cols[e] - co;
        const int64_t i = trans ? c : r, l = trans ? r : c;
        if (i < 0 || i >= M || l < 0 || l >= K) continue;
        const T v = alpha * vals[e];
        const T* x = X + l * sx_l;
        T* out = C + i * sc_i;
        for (int64_t j = lane; j < N; j += 32) atomicAdd(out + j * sc_j, v * x[j * sx_j]);
    }
}

// One thread per entry, walking the columns (use when columns of X are contiguous: consecutive
// entries of a SASO/CSC-ordered operator touch consecutive rows of X).
template <typename T, typename I>
__global__ void __launch_bounds__(256) coo_scatter_cols_kernel(int64_t nnz, const I* __restrict__ rows, const I* __restrict__ cols,
                                                              const T* __restrict__ vals, bool trans, int64_t ro, int64_t co,
                                                              int64_t M, int64_t N, int64_t K, T alpha, const T* __restrict__ X,
                                                              int64_t sx_l, int64_t sx_j, T* __restrict__ C, int64_t sc_i, int64_t sc_j,
                                                              int64_t jchunk) {
    const int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nnz) return;
    const int64_t r = (int64_t) rows[e] - ro, c = (int64_t) cols[e] - co;
    const int64_t i = trans ? c : r, l = trans ? r : c;
    if (i < 0 || i >= M || l < 0 || l >= K) return;
    const T v = alpha * vals[e];
    const int64_t j0 = (int64_t) blockIdx.y * jchunk;
    const int64_t j1 = (j0 + jchunk < N) ? j0 + jchunk : N;
    for (int64_t j = j0; j < j1; ++j) atomicAdd(C + i * sc_i + j * sc_j, v * X[l * sx_l + j * sx_j]);
}

template <typename T, typename I>
int launch_coo_scatter(int64_t nnz, const I* rows, const I* cols, const T* vals, bool trans, int64_t ro, int64_t co,
                       int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                       T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (nnz <= 0 || M <= 0 || N <= 0 || K <= 0 || alpha == (T) 0) return RBX_OK;
    if (sx_j <= sx_l) {
        int64_t blocks = (nnz * 32 + 255) / 256;
        if (blocks > kNumSMs * 32) blocks = kNumSMs * 32;
        coo_scatter_rows_kernel<T, I><<<(unsigned) blocks, 256, 0, stream>>>(nnz, rows, cols, vals, trans, ro, co, M, N, K, alpha,
                                                                           X, sx_l, sx_j, C, sc_i, sc_j);
    } else {
        const int64_t jchunk = 32;
        dim3 grid((unsigned)((nnz + 255) / 256), (unsigned)((N + jchunk - 1) / jchunk));
        if (grid.y > 65535u) return RBX_ERR_UNSUPPORTED;
        coo_scatter_cols_kernel<T, I><<<grid, 256, 0, stream>>>(nnz, rows, cols, vals, trans, ro, co, M, N, K, alpha,
                                                              X, sx_l, sx_j, C, sc_i, sc_j, jchunk);
    }
    count_launch();
    return check_launch("coo_scatter_kernel");
}
#define RBX_INST_COO(T, I)                                                                                                     \
    template int launch_coo_scatter<T, I>(int64_t, const I*, const I*, const T*, bool, int64_t, int64_t, int64_t, int64_t,      \
                                          int64_t, T, const T*, int64_t, int64_t, T*, int64_t, int64_t, cudaStream_t);
RBX_INST_COO(float, int32_t) RBX_INST_COO(float, int64_t) RBX_INST_COO(double, int32_t) RBX_INST_COO(double, int64_t)
#undef RBX_INST_COO

// Sparse data entry (p, q, v) contributes v * Sd[:, l] to C[:, j]; one warp per entry, lanes along i.
template <typename T>
__global__ void __launch_bounds__(256) spdata_scatter_kernel(int64_t nnz, const int64_t* __restrict__ prow, const int64_t* __restrict__ pcol,
                                                            const T* __restrict__ vals, bool trans, int64_t ro, int64_t co,
                                                            int64_t M, int64_t N, int64_t K, T alpha, const T* __restrict__ Sd,
                                                            T* __restrict__ C, int64_t sc_i, int64_t sc_j) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t e = warp; e < nnz; e += nwarps) {
        const int64_t p = prow[e] - ro, q = pcol[e] - co;
        const int64_t l = trans ? q : p, j = trans ? p : q;
        if (l < 0 || l >= K || j < 0 || j >= N) continue;
        const T v = alpha * vals[e];
        const T* s = Sd + l * M;
        T* out = C + j * sc_j;
        for (int64_t i = lane; i < M; i += 32) atomicAdd(out + i * sc_i, v * s[i]);
    }
}

template <typename T>
int launch_spdata_scatter(int64_t nnz, const int64_t* prow, const int64_t* pcol, const T* vals, bool trans, int64_t ro, int64_t co,
                          int64_t M, int64_t N, int64_t K, T alpha, const T* Sd, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (nnz <= 0 || M <= 0 || N <= 0 || K <= 0 || alpha == (T) 0) return RBX_OK;
    int64_t blocks = (nnz * 32 + 255) / 256;
    if (blocks > kNumSMs * 32) blocks = kNumSMs * 32;
    spdata_scatter_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(nnz, prow, pcol, vals, trans, ro, co, M, N, K, alpha, Sd, C, sc_i, sc_j);
    count_launch();
    return check_launch("spdata_scatter_kernel");
}
template int launch_spdata_scatter<float>(int64_t, const int64_t*, const int64_t*, const float*, bool, int64_t, int64_t, int64_t, int64_t,
                                          int64_t, float, const float*, float*, int64_t, int64_t, cudaStream_t);
template int launch_spdata_scatter<double>(int64_t, const int64_t*, const int64_t*, const double*, bool, int64_t, int64_t, int64_t, int64_t,
                                           int64_t, double, const double*, double*, int64_t, int64_t, cudaStream_t);

// dst[l * ldd + j] = src[l + j * lds]  (K x N): column-major data re-laid with contiguous rows, 64x64 tiles through smem
template <typename T>
__global__ void __launch_bounds__(256) transpose_kernel(const T* __restrict__ src, int64_t K, int64_t N, int64_t lds, T* __restrict__ dst, int64_t ldd) {
    __shared__ T tile[64][65];
    const int64_t l0 = (int64_t) blockIdx.x * 64, j0 = (int64_t) blockIdx.y * 64;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    for (int jj = ty; jj < 64; jj += 4) {
        const int64_t l = l0 + tx, j = j0 + jj;
        tile[jj][tx] = (l < K && j < N) ? __ldcs(src + l + j * lds) : (T) 0;
    }
    __syncthreads();
    for (int ll = ty; ll < 64; ll += 4) {
        const int64_t l = l0 + ll, j = j0 + tx;
        if (l < K && j < N) dst[l * ldd + j] = tile[tx][ll];
    }
}
template <typename T>
int launch_transpose(const T* src, int64_t K, int64_t N, int64_t lds, T* dst, int64_t ldd, cudaStream_t stream) {
    if (K <= 0 || N <= 0) return RBX_OK;
    dim3 grid((unsigned)((K + 63) / 64), (unsigned)((N + 63) / 64));
    if (grid.y > 65535u) return RBX_ERR_UNSUPPORTED;
    transpose_kernel<T><<<grid, 256, 0, stream>>>(src, K, N, lds, dst, ldd);
    count_launch();
    return check_launch("transpose_kernel");
}
template int launch_transpose<float>(const float*, int64_t, int64_t, int64_t, float*, int64_t, cudaStream_t);
template int launch_transpose<double>(const double*, int64_t, int64_t, int64_t, double*, int64_t, cudaStream_t);

__global__ void __launch_bounds__(256) expand_ptr_kernel(int64_t n_outer, const int64_t* __restrict__ ptr, int64_t* __restrict__ outer_idx) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t o = warp; o < n_outer; o += nwarps)
        for (int64_t e = ptr[o] + lane; e < ptr[o + 1]; e += 32) outer_idx[e] = o;
}
int launch_expand_ptr(int64_t n_outer, const int64_t* ptr, int64_t* outer_idx, cudaStream_t stream) {
    if (n_outer <= 0) return RBX_OK;
    int64_t blocks = (n_outer * 32 + 255) / 256;
    if (blocks > kNumSMs * 32) blocks = kNumSMs * 32;
    expand_ptr_kernel<<<(unsigned) blocks, 256, 0, stream>>>(n_outer, ptr, outer_idx);
    count_launch();
    return check_launch("expand_ptr_kernel");
}

}  // namespace rbx
