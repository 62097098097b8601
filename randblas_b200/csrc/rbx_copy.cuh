// Re-layout copies that feed the tensor-core sketch kernels (rbx_sketch_dense_tc.cu, rbx_sketch_dense_dmma.cu). Those
// kernels want both operands with unit stride along the contraction axis and 16-byte aligned rows; an operand that is
// laid out the other way round, or whose base / leading dimension is not aligned, is copied chunk by chunk along the
// contraction axis into a k-contiguous workspace first. One extra pass over the operand against a ~16x slower SIMT
// kernel. No reference counterpart: blas::gemm takes any layout (skge.hh:173-215).
#pragma once
#include "rbx_kernels.cuh"

namespace rbx {

// dst[j * pitch + l] = src[l * s_l + j] for a kl x n piece whose rows l are the contiguous ones (unit stride along j):
// 64 x 64 tiles through shared memory, both sides coalesced
template <typename T>
__global__ void __launch_bounds__(256) kmajor_copy_kernel(const T* __restrict__ src, int64_t kl, int64_t n, int64_t s_l,
                                                          T* __restrict__ dst, int64_t pitch) {
    __shared__ T tile[64][65];
    const int64_t l0 = (int64_t) blockIdx.x * 64;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    for (int64_t j0 = (int64_t) blockIdx.y * 64; j0 < n; j0 += (int64_t) gridDim.y * 64) {
        for (int ll = ty; ll < 64; ll += 4) {
            const int64_t l = l0 + ll, j = j0 + tx;
            tile[ll][tx] = (l < kl && j < n) ? __ldcs(src + l * s_l + j) : T(0);
        }
        __syncthreads();
        for (int jj = ty; jj < 64; jj += 4) {
            const int64_t l = l0 + tx, j = j0 + jj;
            if (l < kl && j < n) dst[j * pitch + l] = tile[tx][jj];
        }
        __syncthreads();
    }
}

// dst[j * pitch + l] = src[l + j * s_j]: k-contiguous rows whose base or leading dimension is not 16-byte aligned, repacked
template <typename T>
__global__ void __launch_bounds__(256) repack_kernel(const T* __restrict__ src, int64_t kl, int64_t n, int64_t s_j,
                                                     T* __restrict__ dst, int64_t pitch) {
    for (int64_t j = blockIdx.y; j < n; j += gridDim.y)
        for (int64_t l = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; l < kl; l += (int64_t) gridDim.x * blockDim.x)
            dst[j * pitch + l] = __ldcs(src + l + j * s_j);
}

// dst[j * pitch + l] = src[l * s_l + j * s_j], l < kl, j < n; one of the two strides is 1
template <typename T>
static inline int launch_kmajor_copy(const T* src, int64_t kl, int64_t n, int64_t s_l, int64_t s_j, T* dst, int64_t pitch,
                                     cudaStream_t stream) {
    if (kl <= 0 || n <= 0) return RBX_OK;
    if (s_l == 1) {
        int64_t bx = (kl + 255) / 256; if (bx > 64) bx = 64;
        const int64_t by = n < 65535 ? n : 65535;
        repack_kernel<T><<<dim3((unsigned) bx, (unsigned) by), 256, 0, stream>>>(src, kl, n, s_j, dst, pitch);
    } else {
        if (s_j != 1 || (kl + 63) / 64 > 0x7fffffffll) return RBX_ERR_UNSUPPORTED;
        int64_t by = (n + 63) / 64; if (by > 65535) by = 65535;
        kmajor_copy_kernel<T><<<dim3((unsigned)((kl + 63) / 64), (unsigned) by), 256, 0, stream>>>(src, kl, n, s_l, dst, pitch);
    }
    count_launch();
    return check_launch("kmajor_copy_kernel");
}

}  // namespace rbx
