// K1: dense operator generation on the device.
//
// Replaces dense::fill_dense_submat_impl (dense_skops.hh:96-170, OpenMP over rows) and the layout
// handling of fill_dense (dense_skops.hh:486-532). One work item = one Philox call = four
// consecutive entries of one row of the (row-padded) parent matrix; items are enumerated row by row
// over the window so that consecutive threads write consecutive addresses. The result is a pure
// function of (seed, window) and therefore identical for any grid size (test_denseskop.cc:299-341).
//
// Roofline: 4 (or 8) bytes written per sample, nothing read => HBM-write bound for the uniform
// family; the Gaussian family adds ~30 FP64 operations per sample pair and is FP64-issue bound.
#include "rbx_kernels.cuh"

namespace rbx {

struct FillParams {
    Ctr128 ctr; Key64 key;
    int64_t n_srows, n_scols;   // window size in parent coordinates
    int64_t r0, c0;             // window origin in parent coordinates
    int64_t W;                  // counters per parent row
    int64_t blk0, nblk;         // first 4-block touched in a row, blocks per row
    int64_t rs, cs;             // output strides for parent (row, col)
    int64_t items;              // n_srows * nblk
    int64_t step_i, step_b;     // grid stride decomposed as (rows, blocks)
    int rows_fastest;           // 1: consecutive items walk down a column of blocks (transposed output)
};

template <typename T> __device__ __forceinline__ T cast_scale(float v, T scale, bool do_scale);
template <> __device__ __forceinline__ float cast_scale<float>(float v, float scale, bool do_scale) {
    return do_scale ? __fmul_rn(v, scale) : v;
}
template <> __device__ __forceinline__ double cast_scale<double>(float v, double scale, bool do_scale) {
    return do_scale ? __dmul_rn((double) v, scale) : (double) v;
}

template <typename T> __device__ __forceinline__ void store4(T* p, T a, T b, T c, T d);
template <> __device__ __forceinline__ void store4<float>(float* p, float a, float b, float c, float d) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(a, b, c, d));
}
template <> __device__ __forceinline__ void store4<double>(double* p, double a, double b, double c, double d) {
    __stcs(reinterpret_cast<double2*>(p), make_double2(a, b));
    __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(c, d));
}

#ifndef RBX_FILL_UNROLL
#define RBX_FILL_UNROLL 4      // measured: 1 -> 4 gives +2 % for both families (loop bookkeeping amortised)
#endif
constexpr int kFillUnroll = RBX_FILL_UNROLL;
// the whole-row fill's loop: item k of this thread is counter (lo + k * stride, hi) and float4 dst + 4 * k * stride
template <typename T, char FAMILY, bool SCALE>
__device__ __forceinline__ void linear_loop(T* __restrict__ dst, uint64_t lo, uint32_t c2, uint32_t c3, const Key64 key, int n_it, int64_t stride,
                                            T scale, const double* __restrict__ tab, const double* __restrict__ kl) {
#pragma unroll kFillUnroll
    for (int k = 0; k < n_it; ++k) {
        const float4 v = generate4<FAMILY>(philox4x32_10((uint32_t) lo, (uint32_t)(lo >> 32), c2, c3, key.v[0], key.v[1]), tab, kl);
        store4<T>(dst, cast_scale<T>(v.x, scale, SCALE), cast_scale<T>(v.y, scale, SCALE), cast_scale<T>(v.z, scale, SCALE), cast_scale<T>(v.w, scale, SCALE));
        lo += (uint64_t) stride;
        dst += 4 * stride;
    }
}

// VEC: every block of every row is complete, unit column stride, and 4-element aligned in memory.
// LINEAR (with VEC): the window is whole parent rows stored back to back, so item k is counter k and float4 k -- the
// whole-matrix fill (BASELINE config 3). The general walk's 64-bit (row, block) bookkeeping with its wrap correction is a
// third of the uniform kernel's instructions; this loop has none of it.
template <typename T, char FAMILY, bool VEC, bool LINEAR>
__global__ void __launch_bounds__(256) fill_dense_kernel(T* __restrict__ out, const FillParams p, T scale, bool do_scale) {
    __shared__ double tab[32];
    if (FAMILY == 'G') { load_logf_tab(tab); __syncthreads(); }

    const int64_t tid = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= p.items) return;
    // only the uniform family is ever scaled (sqrt(3), dense_skops.hh:512); for the Gaussian family the flag folds to a constant
    // and the four predicated multiplies leave the loop
    if (FAMILY == 'G') do_scale = false;
    if (LINEAR) {
        double kl[14];
        if (FAMILY == 'G') hoist_bm_consts(kl);
        const Ctr128 base = ctr_add(p.ctr, (uint64_t)(p.r0 * p.W));
        const int64_t stride = (int64_t) gridDim.x * blockDim.x;
        // The 128-bit counter of item `it` is base + it. Unless the low 64 bits wrap inside this launch (they do once in 2^64
        // counters; the general loop below takes that case), the high words are constants and the low words a running 64-bit
        // sum: no carry chain, a 32-bit trip count and a running output pointer. The uniform loop is bound by the integer ALU
        // pipe (20 LOP3 of Philox + bookkeeping, two cycles each), so every add / compare taken out of it shows.
        const uint64_t base_lo = ((uint64_t) base.v[1] << 32) | base.v[0];
        if (base_lo + (uint64_t) p.items >= base_lo) {
            const int n_it = (int)((p.items - tid + stride - 1) / stride);
            if (do_scale) linear_loop<T, FAMILY, true>(out + 4 * tid, base_lo + (uint64_t) tid, base.v[2], base.v[3], p.key, n_it, stride, scale, tab, kl);
            else          linear_loop<T, FAMILY, false>(out + 4 * tid, base_lo + (uint64_t) tid, base.v[2], base.v[3], p.key, n_it, stride, scale, tab, kl);
            return;
        }
        for (int64_t it = tid; it < p.items; it += stride) {
            const float4 v = generate4<FAMILY>(philox4x32_10(ctr_add(base, (uint64_t) it), p.key), tab, kl);
            store4<T>(out + 4 * it, cast_scale<T>(v.x, scale, do_scale), cast_scale<T>(v.y, scale, do_scale),
                      cast_scale<T>(v.z, scale, do_scale), cast_scale<T>(v.w, scale, do_scale));
        }
        return;
    }
    // decompose the first item once, then walk with the precomputed (rows, blocks) stride
    int64_t i, b;
    if (p.rows_fastest) { b = tid / p.n_srows; i = tid - b * p.n_srows; }
    else                { i = tid / p.nblk;    b = tid - i * p.nblk; }
    double kr[14];
    if (FAMILY == 'G') hoist_bm_consts(kr);
    // running counter offset and output pointer, advanced by the (rows, blocks) stride with one wrap correction
    uint64_t coff = (uint64_t)((p.r0 + i) * p.W + p.blk0 + b);
    T* dst = out + i * p.rs + ((p.blk0 + b) * 4 - p.c0) * p.cs;
    const uint64_t coff_step = (uint64_t)(p.step_i * p.W + p.step_b);
    const int64_t dst_step = p.step_i * p.rs + p.step_b * 4 * p.cs;
    const uint64_t coff_wrap = p.rows_fastest ? (uint64_t)(1 - p.n_srows * p.W) : (uint64_t)(p.W - p.nblk);
    const int64_t dst_wrap = p.rows_fastest ? 4 * p.cs - p.n_srows * p.rs : p.rs - p.nblk * 4 * p.cs;
    for (int64_t it = tid; it < p.items; it += (int64_t) gridDim.x * blockDim.x) {
        const Ctr128 c = ctr_add(p.ctr, coff);
        const float4 v = generate4<FAMILY>(philox4x32_10(c, p.key), tab, kr);
        const T e0 = cast_scale<T>(v.x, scale, do_scale), e1 = cast_scale<T>(v.y, scale, do_scale),
                e2 = cast_scale<T>(v.z, scale, do_scale), e3 = cast_scale<T>(v.w, scale, do_scale);
        if (VEC) {
            store4<T>(dst, e0, e1, e2, e3);
        } else {
            const int64_t j0 = (p.blk0 + b) * 4 - p.c0;     // window column of lane 0 (may be negative)
            if (j0 >= 0 && j0 < p.n_scols)         dst[0] = e0;
            if (j0 + 1 >= 0 && j0 + 1 < p.n_scols) dst[p.cs] = e1;
            if (j0 + 2 >= 0 && j0 + 2 < p.n_scols) dst[2 * p.cs] = e2;
            if (j0 + 3 >= 0 && j0 + 3 < p.n_scols) dst[3 * p.cs] = e3;
        }
        coff += coff_step; dst += dst_step;
        if (p.rows_fastest) { i += p.step_i; b += p.step_b; if (i >= p.n_srows) { i -= p.n_srows; b += 1; coff += coff_wrap; dst += dst_wrap; } }
        else                { b += p.step_b; i += p.step_i; if (b >= p.nblk)    { b -= p.nblk;    i += 1; coff += coff_wrap; dst += dst_wrap; } }
    }
}

// Launch: out[(i * rs) + (j * cs)] = parent(r0 + i, c0 + j) for the n_srows x n_scols window.
template <typename T>
int launch_fill_window(char family, int64_t n_cols_parent, T* out, int64_t n_srows, int64_t n_scols, int64_t r0, int64_t c0,
                       const Ctr128& ctr, const Key64& key, int64_t rs, int64_t cs, bool do_scale, cudaStream_t stream) {
    if (n_srows <= 0 || n_scols <= 0) return RBX_OK;
    FillParams p;
    p.ctr = ctr; p.key = key;
    p.n_srows = n_srows; p.n_scols = n_scols; p.r0 = r0; p.c0 = c0;
    p.W = (n_cols_parent + 3) / 4;
    p.blk0 = c0 / 4;
    p.nblk = (c0 + n_scols - 1) / 4 - p.blk0 + 1;
    p.rs = rs; p.cs = cs;
    p.items = n_srows * p.nblk;
    p.rows_fastest = (cs != 1 && rs == 1) ? 1 : 0;
    const bool vec = (cs == 1) && (c0 % 4 == 0) && (n_scols % 4 == 0) && (rs % 4 == 0) &&
                     (reinterpret_cast<uintptr_t>(out) % (4 * sizeof(T)) == 0);
    const int threads = 256;
    // enough CTAs for 8 waves of 8 resident CTAs per SM, but never more than one item per thread needs
    int64_t blocks = (p.items + threads - 1) / threads;
    const int64_t cap = (int64_t) kNumSMs * 8 * 4;
    if (blocks > cap) blocks = cap;
    const int64_t stride = blocks * threads;
    if (p.rows_fastest) { p.step_b = stride / n_srows; p.step_i = stride % n_srows; }
    else                { p.step_i = stride / p.nblk;  p.step_b = stride % p.nblk; }
    const T scale = (T) sqrt(3.0);   // (T)std::sqrt(3): dense_skops.hh:512
    dim3 g((unsigned) blocks), b(threads);
    const bool linear = vec && !p.rows_fastest && c0 == 0 && p.nblk == p.W && n_scols == 4 * p.nblk && rs == n_scols;
#define RBX_FILL_LAUNCH(FAM, V, L) fill_dense_kernel<T, FAM, V, L><<<g, b, 0, stream>>>(out, p, scale, do_scale)
    if (family == 'G') { if (linear) RBX_FILL_LAUNCH('G', true, true); else if (vec) RBX_FILL_LAUNCH('G', true, false); else RBX_FILL_LAUNCH('G', false, false); }
    else               { if (linear) RBX_FILL_LAUNCH('U', true, true); else if (vec) RBX_FILL_LAUNCH('U', true, false); else RBX_FILL_LAUNCH('U', false, false); }
#undef RBX_FILL_LAUNCH
    count_launch();
    return check_launch("fill_dense_kernel");
}

template int launch_fill_window<float>(char, int64_t, float*, int64_t, int64_t, int64_t, int64_t, const Ctr128&, const Key64&,
                                       int64_t, int64_t, bool, cudaStream_t);
template int launch_fill_window<double>(char, int64_t, double*, int64_t, int64_t, int64_t, int64_t, const Ctr128&, const Key64&,
                                        int64_t, int64_t, bool, cudaStream_t);

// ---- raw generator kernels (KAT / transform tests) -------------------------------------------------
__global__ void philox_stream_kernel(Ctr128 ctr, Key64 key, int64_t n, uint4* __restrict__ out) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = philox4x32_10(ctr_add(ctr, (uint64_t) i), key);
}
template <char FAMILY>
__global__ void generate4_stream_kernel(Ctr128 ctr, Key64 key, int64_t n, float4* __restrict__ out) {
    __shared__ double tab[32];
    load_logf_tab(tab); __syncthreads();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = generate4<FAMILY>(philox4x32_10(ctr_add(ctr, (uint64_t) i), key), tab);
}
__global__ void boxmuller_raw_kernel(int64_t n, const uint32_t* __restrict__ u0, const uint32_t* __restrict__ u1,
                                     float* __restrict__ s, float* __restrict__ c) {
    __shared__ double tab[32];
    load_logf_tab(tab); __syncthreads();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float2 r = boxmuller(u0[i], u1[i], tab);
        s[i] = r.x; c[i] = r.y;
    }
}

int launch_philox_stream(const Ctr128& ctr, const Key64& key, int64_t n, uint32_t* out, cudaStream_t stream) {
    if (n <= 0) return RBX_OK;
    philox_stream_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(ctr, key, n, reinterpret_cast<uint4*>(out));
    count_launch();
    return check_launch("philox_stream_kernel");
}
int launch_generate4_stream(char family, const Ctr128& ctr, const Key64& key, int64_t n, float* out, cudaStream_t stream) {
    if (n <= 0) return RBX_OK;
    const unsigned g = (unsigned)((n + 255) / 256);
    if (family == 'G') generate4_stream_kernel<'G'><<<g, 256, 0, stream>>>(ctr, key, n, reinterpret_cast<float4*>(out));
    else               generate4_stream_kernel<'U'><<<g, 256, 0, stream>>>(ctr, key, n, reinterpret_cast<float4*>(out));
    count_launch();
    return check_launch("generate4_stream_kernel");
}
int launch_boxmuller_raw(int64_t n, const uint32_t* u0, const uint32_t* u1, float* s, float* c, cudaStream_t stream) {
    if (n <= 0) return RBX_OK;
    int64_t g = (n + 255) / 256; if (g > kNumSMs * 32) g = kNumSMs * 32;
    boxmuller_raw_kernel<<<(unsigned) g, 256, 0, stream>>>(n, u0, u1, s, c);
    count_launch();
    return check_launch("boxmuller_raw_kernel");
}

}  // namespace rbx
