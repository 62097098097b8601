// K3: sparse operator generation (SASO / LASO index + sign draws) on the device.
//
// Replaces sparse::repeated_fisher_yates (sparse_skops.hh:53-113), which is a serial loop on the CPU.
// The reference runs, for each minor-axis vector i, vec_nnz steps of a Fisher-Yates shuffle on a
// work array of length dim_major that it then restores to the identity (:93-103). Vectors are
// therefore independent: vector i, step j consumes Philox counter seed + i*vec_nnz + j (:78) and only
// words 0 (index) and 1 (sign) of the output. Here one thread owns one vector and keeps the shuffle
// lazily: only displaced positions are remembered, in a per-thread list of at most vec_nnz pairs.
// Bit-exact by construction (integer arithmetic only).
#include "rbx_kernels.cuh"
#include "rbx_fy.cuh"

namespace rbx {

constexpr int kLazyMax = 128;   // longest vec_nnz handled by the per-thread lazy list

template <typename T, typename I, int KMAX>
__global__ void __launch_bounds__(128) fisher_yates_lazy_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t dim_minor,
                                                               I* __restrict__ idxs_major, I* __restrict__ idxs_minor, T* __restrict__ vals) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= dim_minor) return;
    fy_vector<KMAX>(ctr, key, i, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
        const int64_t o = i * vec_nnz + j;
        idxs_major[o] = (I) idx;
        if (vals) vals[o] = negative ? (T) -1.0 : (T) 1.0;
        if (idxs_minor) idxs_minor[o] = (I) i;
    });
}

// Fallback for vec_nnz > kLazyMax: one CTA walks vectors with an explicit work array in global scratch
// (one array per CTA), exactly as the reference does, including the undo pass.
template <typename T, typename I>
__global__ void __launch_bounds__(32) fisher_yates_workarray_kernel(Ctr128 ctr, Key64 key, int64_t vec_nnz, int64_t dim_major, int64_t dim_minor,
                                                                   I* __restrict__ idxs_major, I* __restrict__ idxs_minor, T* __restrict__ vals,
                                                                   int64_t* __restrict__ scratch) {
    int64_t* work = scratch + (int64_t) blockIdx.x * (dim_major + vec_nnz);
    int64_t* pivots = work + dim_major;
    for (int64_t t = threadIdx.x; t < dim_major; t += blockDim.x) work[t] = t;
    __syncwarp();
    if (threadIdx.x != 0) return;
    for (int64_t i = blockIdx.x; i < dim_minor; i += gridDim.x) {
        for (int64_t j = 0; j < vec_nnz; ++j) {
            const uint4 rv = philox4x32_10(ctr_add(ctr, (uint64_t)(i * vec_nnz + j)), key);
            const int64_t ell = j + (int64_t)((uint64_t) rv.x % (uint64_t)(dim_major - j));
            pivots[j] = ell;
            const int64_t swap = work[ell];
            work[ell] = work[j];
            work[j] = swap;
            const int64_t o = i * vec_nnz + j;
            idxs_major[o] = (I) swap;
            if (vals) vals[o] = (rv.y % 2 == 0) ? (T) 1.0 : (T) -1.0;
            if (idxs_minor) idxs_minor[o] = (I) i;
        }
        for (int64_t j = vec_nnz - 1; j >= 0; --j) {   // restore the identity (sparse_skops.hh:93-103)
            const int64_t ell = pivots[j];
            const int64_t swap = work[ell];
            work[ell] = work[j];
            work[j] = swap;
        }
    }
}

template <typename T, typename I>
int launch_fisher_yates(const Ctr128& ctr, const Key64& key, int64_t vec_nnz, int64_t dim_major, int64_t dim_minor,
                        I* idxs_major, I* idxs_minor, T* vals, cudaStream_t stream) {
    if (vec_nnz > dim_major) return RBX_ERR_ARG;     // randblas_error_if(vec_nnz > dim_major), sparse_skops.hh:63
    if (vec_nnz <= 0 || dim_minor <= 0) return RBX_OK;
    if (vec_nnz <= kLazyMax) {
        const unsigned g = (unsigned)((dim_minor + 127) / 128);
        if (vec_nnz <= 8)       fisher_yates_lazy_kernel<T, I, 8><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        else if (vec_nnz <= 32) fisher_yates_lazy_kernel<T, I, 32><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        else                    fisher_yates_lazy_kernel<T, I, kLazyMax><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        count_launch();
        return check_launch("fisher_yates_lazy_kernel");
    }
    int64_t ctas = dim_minor < kNumSMs * 4 ? dim_minor : kNumSMs * 4;
    // one work array of (dim_major + vec_nnz) int64 per CTA: the CTA count follows the memory that is free (at most 1 GiB)
    const size_t per_cta = sizeof(int64_t) * (size_t)(dim_major + vec_nnz);
    const int64_t fit = (int64_t)(ws_budget((size_t) 1 << 30) / per_cta);
    if (fit < 1) return RBX_ERR_ALLOC;
    if (ctas > fit) ctas = fit;
    int64_t* scratch = nullptr;
    if (ws_alloc(&scratch, sizeof(int64_t) * (size_t)(ctas * (dim_major + vec_nnz)), stream) != cudaSuccess) {
        cudaGetLastError();
        return RBX_ERR_ALLOC;
    }
    fisher_yates_workarray_kernel<T, I><<<(unsigned) ctas, 32, 0, stream>>>(ctr, key, vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals, scratch);
    count_launch();
    int rc = check_launch("fisher_yates_workarray_kernel");
    cudaFreeAsync(scratch, stream);
    return rc;
}

#define RBX_INST_FY(T, I) \
    template int launch_fisher_yates<T, I>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, I*, I*, T*, cudaStream_t);
RBX_INST_FY(float, int32_t) RBX_INST_FY(float, int64_t) RBX_INST_FY(double, int32_t) RBX_INST_FY(double, int64_t)
#undef RBX_INST_FY

}  // namespace rbx
