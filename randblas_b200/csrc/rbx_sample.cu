// Index sampling utilities on the device: util::sample_indices_iid_uniform, util::sample_indices_iid and
// util::weights_to_cdf (util.hh:257-334). Same Philox + uneg11 stream as the reference, bit for bit:
// sample i uses lane i % 4 of uneg11(philox(ctr + 1 + i / 4)) (the reference increments the counter before its first
// draw), mapped to [0, 1] as (x + 1) / 2 in TF precision, then index = trunc(n * u) or lower_bound(cdf, u).
// One thread per counter (four samples); the cdf form does four binary searches over the device-resident cdf.
// weights_to_cdf keeps the reference's sequential running sum (the order defines the bits): one warp loads 32 weights
// at a time and replays them in order through shuffles.
#include "rbx_kernels.cuh"

namespace rbx {

__global__ void __launch_bounds__(256) sample_uniform_kernel(int64_t n, int64_t k, int64_t* __restrict__ samples, Ctr128 ctr, Key64 key) {
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c * 4 >= k) return;
    const uint4 r = philox4x32_10(ctr_add(ctr, (uint64_t)(1 + c)), key);
    const uint32_t rv[4] = { r.x, r.y, r.z, r.w };
    const double dn = (double) n;        // `(int64_t) dN * u` in the reference is ((int64_t) dN) * u: a double product
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int64_t i = c * 4 + q;
        if (i < k) {
            const double u = __ddiv_rn(__dadd_rn((double) uneg11(rv[q]), 1.0), 2.0);
            samples[i] = (int64_t) __dmul_rn(dn, u);
        }
    }
}

template <typename T> __device__ __forceinline__ T to_unit(float x);
template <> __device__ __forceinline__ float to_unit<float>(float x) { return __fdiv_rn(__fadd_rn(x, 1.0f), 2.0f); }
template <> __device__ __forceinline__ double to_unit<double>(float x) { return __ddiv_rn(__dadd_rn((double) x, 1.0), 2.0); }

template <typename T>
__global__ void __launch_bounds__(256) sample_cdf_kernel(int64_t n, const T* __restrict__ cdf, int64_t k, int64_t* __restrict__ samples,
                                                        Ctr128 ctr, Key64 key) {
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c * 4 >= k) return;
    const uint4 r = philox4x32_10(ctr_add(ctr, (uint64_t)(1 + c)), key);
    const uint32_t rv[4] = { r.x, r.y, r.z, r.w };
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int64_t i = c * 4 + q;
        if (i < k) {
            const T u = to_unit<T>(uneg11(rv[q]));
            int64_t lo = 0, hi = n;                       // std::lower_bound: first index with cdf[idx] >= u
            while (lo < hi) { const int64_t mid = lo + ((hi - lo) >> 1); if (cdf[mid] < u) lo = mid + 1; else hi = mid; }
            samples[i] = lo;
        }
    }
}

// status[0] |= 1: a weight below the threshold (or NaN); status[0] |= 2: total too small. sum_out[0] = the total.
template <typename T>
__global__ void __launch_bounds__(32) weights_running_sum_kernel(int64_t n, T* __restrict__ w, T error_if_below, int* __restrict__ status,
                                                                T* __restrict__ sum_out) {
    const int lane = threadIdx.x;
    T sum = (T) 0;
    int bad = 0;
    for (int64_t base = 0; base < n; base += 32) {
        const int64_t i = base + lane;
        T val = (i < n) ? w[i] : (T) 0;
        if (i < n && !(val >= error_if_below)) bad = 1;
        val = val > (T) 0 ? val : (T) 0;                  // std::max(val, 0)
        T mine = (T) 0;
        const int cnt = (n - base < 32) ? (int)(n - base) : 32;
        for (int q = 0; q < cnt; ++q) {
            sum += __shfl_sync(0xffffffffu, val, q);      // every lane replays the sequential sum
            if (q == lane) mine = sum;
        }
        if (i < n) w[i] = mine;
    }
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) {
        const T eps = sizeof(T) == 4 ? (T) 0x1p-23 : (T) 0x1p-52;
        int st = bad ? 1 : 0;
        if (!(sum >= ((T) sqrt((double) n)) * eps)) st |= 2;
        status[0] = st;
        sum_out[0] = sum;
    }
}
template <typename T>
__global__ void __launch_bounds__(256) weights_scale_kernel(int64_t n, T* __restrict__ w, const int* __restrict__ status, const T* __restrict__ sum) {
    if (status[0] != 0) return;                           // the reference throws before scaling
    const T inv = ((T) 1.0) / sum[0];
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) w[i] = inv * w[i];
}

int launch_sample_uniform(int64_t n, int64_t k, int64_t* samples, const Ctr128& ctr, const Key64& key, cudaStream_t stream) {
    if (k <= 0) return RBX_OK;
    const int64_t nc = (k + 3) / 4;
    sample_uniform_kernel<<<(unsigned)((nc + 255) / 256), 256, 0, stream>>>(n, k, samples, ctr, key);
    count_launch();
    return check_launch("sample_uniform_kernel");
}
template <typename T>
int launch_sample_cdf(int64_t n, const T* cdf, int64_t k, int64_t* samples, const Ctr128& ctr, const Key64& key, cudaStream_t stream) {
    if (k <= 0) return RBX_OK;
    const int64_t nc = (k + 3) / 4;
    sample_cdf_kernel<T><<<(unsigned)((nc + 255) / 256), 256, 0, stream>>>(n, cdf, k, samples, ctr, key);
    count_launch();
    return check_launch("sample_cdf_kernel");
}
// returns RBX_ERR_ARG where the reference throws (synchronises the stream to read the status word)
template <typename T>
int launch_weights_to_cdf(int64_t n, T* w, T error_if_below, cudaStream_t stream) {
    int* status = nullptr; T* sum = nullptr;
    if (ws_alloc(&status, sizeof(int), stream) != cudaSuccess || ws_alloc(&sum, sizeof(T), stream) != cudaSuccess) {
        cudaGetLastError(); if (status) cudaFreeAsync(status, stream); return RBX_ERR_ALLOC;
    }
    weights_running_sum_kernel<T><<<1, 32, 0, stream>>>(n, w, error_if_below, status, sum);
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    if (blocks < 1) blocks = 1;
    weights_scale_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(n, w, status, sum);
    count_launch(); count_launch();
    int rc = check_launch("weights_to_cdf kernels");
    int st = 0;
    if (rc == RBX_OK && (cudaMemcpyAsync(&st, status, sizeof(int), cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
                         cudaStreamSynchronize(stream) != cudaSuccess)) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    cudaFreeAsync(status, stream); cudaFreeAsync(sum, stream);
    if (rc == RBX_OK && st != 0) rc = RBX_ERR_ARG;
    return rc;
}
template int launch_sample_cdf<float>(int64_t, const float*, int64_t, int64_t*, const Ctr128&, const Key64&, cudaStream_t);
template int launch_sample_cdf<double>(int64_t, const double*, int64_t, int64_t*, const Ctr128&, const Key64&, cudaStream_t);
template int launch_weights_to_cdf<float>(int64_t, float*, float, cudaStream_t);
template int launch_weights_to_cdf<double>(int64_t, double*, double, cudaStream_t);

}  // namespace rbx
