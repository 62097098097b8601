// Small helper kernels shared by the sparse paths: the safe_scal of an output view (util.hh:51-59), the expansion of a CSR /
// CSC pointer array into one outer index per nonzero (the device-side step of viewing compressed data as COO triplets,
// spmm_dispatch.hh:48-200) and util::require_symmetric (util.hh:165-188). The products themselves live in rbx_saso.cu (SASO
// streaming kernel), rbx_spmm_gather.cu (every other sparse operator, deterministic gather) and rbx_spdata.cu (sparse data).
// (Round 1's atomicAdd scatter lived here; it was replaced by the gather kernels and removed.)
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
