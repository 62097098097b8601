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

// util::require_symmetric (util.hh:165-188) on the device: |A_ij - A_ji| <= (|A_ij| + |A_ji| + 1) * tol for all i < j.
// One thread per (i, j) pair of the strict upper triangle, tiles of 32 x 32 read through both orientations; *violated is
// raised (never cleared) by any thread that finds a pair out of tolerance. NaNs never violate, as in the reference.
template <typename T>
__global__ void __launch_bounds__(256) require_symmetric_kernel(const T* __restrict__ A, int64_t n, int64_t s_row, int64_t s_col, T tol, int* violated) {
    const int64_t tj = blockIdx.x, ti = blockIdx.y;
    if (ti > tj) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int64_t i = ti * 32 + ty + 8 * r, j = tj * 32 + tx;
        if (i < n && j < n && i < j) {
            const T Aij = A[i * s_row + j * s_col], Aji = A[j * s_row + i * s_col];
            const T viol = fabs(Aij - Aji), rel_tol = (fabs(Aij) + fabs(Aji) + (T) 1) * tol;
            if (viol > rel_tol) *violated = 1;
        }
    }
}
template <typename T>
int launch_require_symmetric(const T* A, int64_t n, int64_t s_row, int64_t s_col, T tol, int* violated_dev, cudaStream_t stream) {
    if (n <= 1) return RBX_OK;
    const int64_t nt = (n + 31) / 32;
    if (nt > 65535) return RBX_ERR_UNSUPPORTED;
    require_symmetric_kernel<T><<<dim3((unsigned) nt, (unsigned) nt), 256, 0, stream>>>(A, n, s_row, s_col, tol, violated_dev);
    count_launch();
    return check_launch("require_symmetric_kernel");
}
template int launch_require_symmetric<float>(const float*, int64_t, int64_t, int64_t, float, int*, cudaStream_t);
template int launch_require_symmetric<double>(const double*, int64_t, int64_t, int64_t, double, int*, cudaStream_t);

}  // namespace rbx
