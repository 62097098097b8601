// K5: dense operator x sparse data (sketch_sparse; lsksp3 / rsksp3, sksp.hh:147-345).
//
// The reference materialises S and calls right_spmm / left_spmm, whose kernels (csc_spmm_impl.hh:90-213,
// csr_spmm_impl.hh:69-156, coo_spmm_impl.hh:79-162) make one pass over all nonzeros per row of S. Here the product
// C (M x N) = alpha * Sop (M x K) * op(A) (K x N, sparse) + beta * C is output-stationary: one warp owns 512 rows of one
// column j of C in registers and, for every nonzero (l, v) of column j of op(A), adds v * Sop[:, l]. Nothing is
// accumulated with atomics and the summation order is fixed (entries of a column in increasing l), so the result is
// bitwise reproducible from run to run.
//   * Sop is generated once per slab of the contraction axis by K1 into a FLOAT array, i contiguous (the generator's
//     values are floats by construction, dense_skops.hh:128-131; the uniform sqrt(3) factor is applied in T precision
//     on load exactly as fill_dense does), which halves the gather traffic for T = double.
//   * The nonzeros must be grouped by j. CSC data (NoTrans) and CSR data (Trans) already are; CSR NoTrans, CSC Trans and
//     COO are regrouped on the device: count per column, exclusive scan, scatter, then a rank sort inside each column
//     segment (conversions.hh:43-159 and the std::sort of coo_matrix.hh:267-325 are the host-side counterparts).
// Roofline: HBM. Each nonzero gathers one column of Sop (4 * M bytes); with random sparsity these are L2 misses, so the
// kernel's real floor is nnz * 4 * M bytes of DRAM reads (205 GB at C4), far above the compulsory bytes of SURVEY 8(d).
#include "rbx_kernels.cuh"

namespace rbx {

// ---- regrouping: per-nonzero (p, q, v) -> column segments of op(A) ----------------------------------------------------
__device__ __forceinline__ bool map_entry(int64_t p, int64_t q, bool trans, int64_t ro, int64_t co, int64_t K, int64_t N,
                                          int64_t* l, int64_t* j) {
    p -= ro; q -= co;
    *l = trans ? q : p; *j = trans ? p : q;
    return *l >= 0 && *l < K && *j >= 0 && *j < N;
}

// Segments are keyed by j * nlc + (l >> lc_shift): nlc == 1 groups by column j alone (sketch_sparse, conversions); nlc > 1
// cuts every column into sub-segments by blocks of 2^lc_shift contraction indices, which keeps the segments that the rank
// sort sees short however many entries a column has (the general sparse-operator path, rbx_spmm_gather.cu). Sub-segments
// of one column are adjacent and in increasing l-block order, so a column's entries still come out sorted by l.
__global__ void __launch_bounds__(256) spd_count_kernel(int64_t nnz, const int64_t* __restrict__ prow, const int64_t* __restrict__ pcol,
                                                       bool trans, int64_t ro, int64_t co, int64_t K, int64_t N, int lc_shift, int64_t nlc,
                                                       unsigned long long* __restrict__ cnt) {
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t) gridDim.x * blockDim.x) {
        int64_t l, j;
        if (map_entry(prow[e], pcol[e], trans, ro, co, K, N, &l, &j)) atomicAdd(cnt + (j * nlc + (l >> lc_shift)), 1ull);
    }
}

// exclusive scan of n counters in three phases (1024 per block)
__global__ void __launch_bounds__(256) spd_scan_partial_kernel(int64_t n, const unsigned long long* __restrict__ in,
                                                              unsigned long long* __restrict__ block_sums) {
    __shared__ unsigned long long red[8];
    const int64_t base = (int64_t) blockIdx.x * 1024 + threadIdx.x * 4;
    unsigned long long s = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) if (base + q < n) s += in[base + q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long t = 0; for (int w = 0; w < 8; ++w) t += red[w]; block_sums[blockIdx.x] = t; }
}
__global__ void spd_scan_blocks_kernel(int64_t nb, unsigned long long* __restrict__ block_sums) {   // one thread: nb <= 2^21
    unsigned long long run = 0;
    for (int64_t b = 0; b < nb; ++b) { const unsigned long long v = block_sums[b]; block_sums[b] = run; run += v; }
}
__global__ void __launch_bounds__(256) spd_scan_final_kernel(int64_t n, const unsigned long long* __restrict__ in,
                                                            const unsigned long long* __restrict__ block_sums,
                                                            int64_t* __restrict__ ptr, unsigned long long* __restrict__ cursor) {
    __shared__ unsigned long long wsum[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base = (int64_t) blockIdx.x * 1024 + threadIdx.x * 4;
    unsigned long long v[4], s = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[q] = (base + q < n) ? in[base + q] : 0ull; s += v[q]; }
    unsigned long long inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    unsigned long long off = block_sums[blockIdx.x];
    for (int w = 0; w < warp; ++w) off += wsum[w];
    off += inc - s;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (base + q < n) { ptr[base + q] = (int64_t) off; cursor[base + q] = off; }
        off += v[q];
    }
    if (base <= n && n < base + 4) ptr[n] = (int64_t) off;        // total, written by the thread whose range ends at n
}

// one 16-byte record per regrouped entry, so that the scattered write touches one sector
template <typename T> struct __align__(16) SpdRec { int64_t l; T v; };

template <typename T>
__global__ void __launch_bounds__(256) spd_scatter_kernel(int64_t nnz, const int64_t* __restrict__ prow, const int64_t* __restrict__ pcol,
                                                         const T* __restrict__ vals, bool trans, int64_t ro, int64_t co, int64_t K, int64_t N,
                                                         int lc_shift, int64_t nlc,
                                                         unsigned long long* __restrict__ cursor, SpdRec<T>* __restrict__ out) {
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (int64_t) gridDim.x * blockDim.x) {
        int64_t l, j;
        if (map_entry(prow[e], pcol[e], trans, ro, co, K, N, &l, &j)) {
            const unsigned long long pos = atomicAdd(cursor + (j * nlc + (l >> lc_shift)), 1ull);
            SpdRec<T> r; r.l = l; r.v = vals[e];
            out[pos] = r;
        }
    }
}

template <typename T> struct BitsOf;
template <> struct BitsOf<float>  { __device__ static unsigned long long get(float v)  { return __float_as_uint(v); } };
template <> struct BitsOf<double> { __device__ static unsigned long long get(double v) { return (unsigned long long) __double_as_longlong(v); } };

// One warp per column segment: rank sort by (l, value bits, position). The scatter above places entries in arrival
// order, which varies from run to run; this pass makes the segment - and with it the summation order - canonical.
// Segments of up to kSortSmem entries are ranked out of shared memory (broadcast reads); longer ones out of global
// memory; segments longer than kSortCap are copied as they are (quadratic work; documented limit).
constexpr int64_t kSortCap = 8192;
constexpr int kSortSmem = 256;
template <typename T>
__global__ void __launch_bounds__(256) spd_segment_sort_kernel(int64_t N, const int64_t* __restrict__ ptr, const SpdRec<T>* __restrict__ in,
                                                              int64_t* __restrict__ out_l, T* __restrict__ out_v, bool small_keys) {
    __shared__ int keys[8][kSortSmem];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t j = warp; j < N; j += nwarps) {
        const int64_t beg = ptr[j], len = ptr[j + 1] - beg;
        const bool in_smem = len <= kSortSmem && small_keys;
        __syncwarp();
        if (in_smem) for (int64_t x = lane; x < len; x += 32) keys[wib][x] = (int) in[beg + x].l;
        __syncwarp();
        for (int64_t x = lane; x < len; x += 32) {
            const SpdRec<T> rx = in[beg + x];
            int64_t rank = x;
            if (len <= kSortCap) {
                int less = 0, equal = 0;
                if (in_smem) {
                    const int kx = (int) rx.l, n = (int) len;
#pragma unroll 4
                    for (int y = 0; y < n; ++y) { const int ky = keys[wib][y]; less += (ky < kx); equal += (ky == kx); }
                } else {
                    for (int64_t y = 0; y < len; ++y) { const int64_t ly = in[beg + y].l; less += (ly < rx.l); equal += (ly == rx.l); }
                }
                rank = less;
                if (equal > 1) {                   // duplicates of (l, j): order them by value bits, then by position
                    const unsigned long long bx = BitsOf<T>::get(rx.v);
                    for (int64_t y = 0; y < len; ++y) {
                        const SpdRec<T> ry = in[beg + y];
                        if (ry.l == rx.l) { const unsigned long long by = BitsOf<T>::get(ry.v); rank += (by < bx || (by == bx && y < x)) ? 1 : 0; }
                    }
                }
            }
            out_l[beg + rank] = rx.l; out_v[beg + rank] = rx.v;
        }
    }
}

// ---- the gather kernel ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ T spd_value(float s, bool uniform);
template <> __device__ __forceinline__ float spd_value<float>(float s, bool uniform) {
    return uniform ? __fmul_rn(s, (float) 1.7320508075688772) : s;
}
template <> __device__ __forceinline__ double spd_value<double>(float s, bool uniform) {
    return uniform ? __dmul_rn((double) s, 1.7320508075688772) : (double) s;
}

struct SpdArgs {
    int64_t N; const int64_t* ptr; const int64_t* lidx;   // column segments of op(A): lidx / vals [ptr[j], ptr[j+1])
    int64_t l0, kk;                                        // contraction slab held in Sd
    int64_t M, Mp;                                         // rows of the operator window, slab pitch (multiple of 4)
    const float* Sd;                                       // Sd[(l - l0) * Mp + i]
    int64_t sc_i, sc_j;
    int uniform; int first;                                // first slab applies beta, later slabs accumulate
};

// grid.y tiles M in chunks of 512 rows; each warp walks columns j (grid-stride) and keeps 16 accumulators per lane:
// rows i = chunk + q * 128 + lane * 4 + (0..3), q = 0..3, so every gather is one 16-byte load per lane and q.
template <typename T, bool UNI>
__global__ void __launch_bounds__(256, 2) spdata_gather_kernel(const SpdArgs a, const T* __restrict__ vals, T alpha, T beta, T* __restrict__ C) {
    constexpr int U = sizeof(T) == 8 ? 2 : 4;                      // entries in flight per warp (register budget: 2 CTAs / SM)
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const int64_t ib = (int64_t) blockIdx.y * 512 + lane * 4;
    bool live[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) live[q] = ib + q * 128 < a.M;      // Mp is padded to 4, so a live float4 is always in bounds
    for (int64_t j = warp; j < a.N; j += nwarps) {
        const int64_t beg = a.ptr[j], end = a.ptr[j + 1];
        T acc[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[q][c] = (T) 0;
        for (int64_t e0 = beg; e0 < end; e0 += 32) {
            int64_t my_l = -1; T my_v = (T) 0;
            if (e0 + lane < end) {
                my_l = a.lidx[e0 + lane] - a.l0; my_v = vals[e0 + lane];
                if (my_l < 0 || my_l >= a.kk) { my_l = 0; my_v = (T) 0; }     // outside this slab: contributes v = 0
            }
            const int cnt = (end - e0 < 32) ? (int)(end - e0) : 32;
            for (int u0 = 0; u0 < cnt; u0 += U) {
                float4 s[U][4]; T v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    int64_t l = __shfl_sync(0xffffffffu, my_l, (u0 + u) & 31);
                    v[u] = __shfl_sync(0xffffffffu, my_v, (u0 + u) & 31);
                    if (u0 + u >= cnt) { l = 0; v[u] = (T) 0; }
                    if (l < 0) l = 0;
                    const float* col = a.Sd + l * a.Mp + ib;
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        s[u][q] = live[q] ? __ldg(reinterpret_cast<const float4*>(col + q * 128)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        acc[q][0] = fma(v[u], spd_value<T>(s[u][q].x, UNI), acc[q][0]);
                        acc[q][1] = fma(v[u], spd_value<T>(s[u][q].y, UNI), acc[q][1]);
                        acc[q][2] = fma(v[u], spd_value<T>(s[u][q].z, UNI), acc[q][2]);
                        acc[q][3] = fma(v[u], spd_value<T>(s[u][q].w, UNI), acc[q][3]);
                    }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int64_t i = ib + q * 128 + c;
                if (i < a.M) {
                    T* p = C + i * a.sc_i + j * a.sc_j;
                    const T r = alpha * acc[q][c];
                    if (a.first) *p = (beta == (T) 0) ? r : r + beta * (*p);      // beta == 0: C is not read
                    else         *p += r;
                }
            }
    }
}

template <typename T>
int launch_spdata_gather(int64_t N, const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t l0, int64_t kk, int64_t M, int64_t Mp,
                         const float* Sd, bool uniform, bool first, T alpha, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (N <= 0 || M <= 0) return RBX_OK;
    SpdArgs a;
    a.N = N; a.ptr = ptr; a.lidx = lidx; a.l0 = l0; a.kk = kk; a.M = M; a.Mp = Mp; a.Sd = Sd; a.sc_i = sc_i; a.sc_j = sc_j;
    a.uniform = uniform ? 1 : 0; a.first = first ? 1 : 0;
    const int64_t ychunks = (M + 511) / 512;
    if (ychunks > 65535) return RBX_ERR_UNSUPPORTED;
    int64_t bx = (N + 7) / 8;                                   // 8 warps per CTA, one column per warp and pass
    const int64_t cap = (int64_t) kNumSMs * 8 * 4;              // 8 resident CTAs per SM, 4 waves
    if (bx > cap) bx = cap;
    dim3 grid((unsigned) bx, (unsigned) ychunks);
    if (uniform) spdata_gather_kernel<T, true><<<grid, 256, 0, stream>>>(a, vals, alpha, beta, C);
    else         spdata_gather_kernel<T, false><<<grid, 256, 0, stream>>>(a, vals, alpha, beta, C);
    count_launch();
    return check_launch("spdata_gather_kernel");
}

// ---- the gather kernel for an L2-resident operator slab (d-blocked passes) ----------------------------------------------
// When the whole operator window (4 * M * K bytes as floats) is far larger than L2, every nonzero's gather of M floats is an
// HBM miss (C4 round 1: 198 GB of DRAM reads against 5.7 GB of compulsory bytes). Here the OUTPUT rows are blocked instead:
// one pass handles DC rows of C (DC = 16: a slab of 16 x K floats = 64 MB at C4), generates just those rows of the operator
// (l-major, DC floats = 64 bytes per contraction index) and re-streams the nonzeros of op(A); the slab stays in L2 across
// the pass (evict_last policy), the nonzeros are read with evict_first, C is written once (streaming stores). DRAM traffic
// becomes (M / DC) * nnz * 16 + slab + C instead of nnz * 4 * M.
// Mapping: DC / 4 lanes per nonzero (one float4 of operator rows each), 32 / (DC / 4) nonzeros per warp step, U steps in
// flight; the per-group partial sums are folded by a fixed butterfly => bitwise reproducible.
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
    unsigned long long p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ unsigned long long l2_policy_evict_first() {
    unsigned long long p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ float4 ldg_f4_hint(const float* p, unsigned long long pol) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ unsigned long long ldg_u64_hint(const void* p, unsigned long long pol) {
    unsigned long long v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.b64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
    return v;
}
template <typename T> __device__ __forceinline__ T ldg_val_hint(const T* p, unsigned long long pol);
template <> __device__ __forceinline__ double ldg_val_hint<double>(const double* p, unsigned long long pol) { return __longlong_as_double((long long) ldg_u64_hint(p, pol)); }
template <> __device__ __forceinline__ float ldg_val_hint<float>(const float* p, unsigned long long pol) {
    float v; asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol)); return v;
}

struct SpdNarrowArgs {
    int64_t N; const int64_t* ptr; const int64_t* lidx;
    int64_t K;
    int rows;                                              // live rows of this pass (<= DC)
    const float* Sd;                                       // Sd[l * DC + i]
    int64_t sc_i, sc_j;
};

template <typename T, int DC, bool UNI>
__global__ void __launch_bounds__(256) spdata_gather_narrow_kernel(const SpdNarrowArgs a, const T* __restrict__ vals, T alpha, T beta, T* __restrict__ C) {
    constexpr int LPE = DC / 4, EPW = 32 / LPE, U = 4;
    const int lane = threadIdx.x & 31, g = lane / LPE, sub = lane % LPE;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const unsigned long long keep = l2_policy_evict_last(), stream_once = l2_policy_evict_first();
    const float* sd = a.Sd + sub * 4;
    for (int64_t j = warp; j < a.N; j += nwarps) {
        const int64_t beg = a.ptr[j], end = a.ptr[j + 1];
        T acc[4] = { (T) 0, (T) 0, (T) 0, (T) 0 };
        for (int64_t e0 = beg + g; e0 < end; e0 += EPW * U) {
            int64_t l[U]; T v[U]; float4 s[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t e = e0 + (int64_t) u * EPW;
                const bool in = e < end;
                l[u] = in ? (int64_t) ldg_u64_hint(a.lidx + e, stream_once) : 0;
                v[u] = in ? ldg_val_hint<T>(vals + e, stream_once) : (T) 0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) s[u] = ldg_f4_hint(sd + l[u] * DC, keep);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                acc[0] = fma(v[u], spd_value<T>(s[u].x, UNI), acc[0]);
                acc[1] = fma(v[u], spd_value<T>(s[u].y, UNI), acc[1]);
                acc[2] = fma(v[u], spd_value<T>(s[u].z, UNI), acc[2]);
                acc[3] = fma(v[u], spd_value<T>(s[u].w, UNI), acc[3]);
            }
        }
#pragma unroll
        for (int o = LPE; o < 32; o <<= 1)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], o);
        if (g == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int i = sub * 4 + q;
                if (i < a.rows) {
                    T* p = C + i * a.sc_i + j * a.sc_j;
                    const T r = alpha * acc[q];
                    *p = (beta == (T) 0) ? r : r + beta * (*p);      // beta == 0: C is not read
                }
            }
        }
    }
}

template <typename T>
int launch_spdata_gather_narrow(int64_t N, const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t K, int rows, int dc, const float* Sd,
                                bool uniform, T alpha, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    if (N <= 0 || rows <= 0) return RBX_OK;
    if (dc != 16 && dc != 8 && dc != 32) return RBX_ERR_UNSUPPORTED;
    SpdNarrowArgs a;
    a.N = N; a.ptr = ptr; a.lidx = lidx; a.K = K; a.rows = rows; a.Sd = Sd; a.sc_i = sc_i; a.sc_j = sc_j;
    int64_t bx = (N + 7) / 8;
    const int64_t cap = (int64_t) kNumSMs * 8 * 2;
    if (bx > cap) bx = cap;
#define RBX_NARROW(DCV)                                                                                                          \
    if (dc == DCV) {                                                                                                             \
        if (uniform) spdata_gather_narrow_kernel<T, DCV, true><<<(unsigned) bx, 256, 0, stream>>>(a, vals, alpha, beta, C);        \
        else         spdata_gather_narrow_kernel<T, DCV, false><<<(unsigned) bx, 256, 0, stream>>>(a, vals, alpha, beta, C);       \
    }
    RBX_NARROW(8) RBX_NARROW(16) RBX_NARROW(32)
#undef RBX_NARROW
    count_launch();
    return check_launch("spdata_gather_narrow_kernel");
}
template int launch_spdata_gather_narrow<float>(int64_t, const int64_t*, const int64_t*, const float*, int64_t, int, int, const float*, bool, float, float,
                                                float*, int64_t, int64_t, cudaStream_t);
template int launch_spdata_gather_narrow<double>(int64_t, const int64_t*, const int64_t*, const double*, int64_t, int, int, const float*, bool, double, double,
                                                 double*, int64_t, int64_t, cudaStream_t);

template <typename T>
int spd_group_by_column(int64_t nnz, const int64_t* prow, const int64_t* pcol, const T* vals, bool trans, int64_t ro, int64_t co,
                        int64_t K, int64_t N_cols, int64_t** ptr_out, int64_t** lidx_out, T** vals_out, cudaStream_t stream,
                        int lc_shift, int64_t nlc) {
    *ptr_out = nullptr; *lidx_out = nullptr; *vals_out = nullptr;
    if (nlc < 1) { nlc = 1; lc_shift = 62; }
    if (nlc == 1) lc_shift = 62;                       // l >> 62 == 0 for every valid l
    const int64_t N = N_cols * nlc;                    // number of segments
    const int64_t nb = (N + 1 + 1023) / 1024;
    unsigned long long *cnt = nullptr, *cursor = nullptr, *bsum = nullptr;
    int64_t *ptr = nullptr, *out_l = nullptr; T* out_v = nullptr; SpdRec<T>* tmp = nullptr;
    const size_t nz = (size_t)(nnz > 0 ? nnz : 1);
    bool ok = ws_alloc(&cnt, 8 * (size_t)(N + 1), stream) == cudaSuccess && ws_alloc(&cursor, 8 * (size_t)(N + 1), stream) == cudaSuccess &&
              ws_alloc(&bsum, 8 * (size_t) nb, stream) == cudaSuccess && ws_alloc(&ptr, 8 * (size_t)(N + 1), stream) == cudaSuccess &&
              ws_alloc(&tmp, sizeof(SpdRec<T>) * nz, stream) == cudaSuccess && ws_alloc(&out_l, 8 * nz, stream) == cudaSuccess &&
              ws_alloc(&out_v, sizeof(T) * nz, stream) == cudaSuccess;
    int rc = ok ? RBX_OK : RBX_ERR_ALLOC;
    if (!ok) cudaGetLastError();
    if (rc == RBX_OK && cudaMemsetAsync(cnt, 0, 8 * (size_t)(N + 1), stream) != cudaSuccess) rc = RBX_ERR_CUDA;
    if (rc == RBX_OK) {
        int64_t blocks = (nnz + 255) / 256;
        if (blocks > kNumSMs * 32) blocks = kNumSMs * 32;
        if (blocks < 1) blocks = 1;
        spd_count_kernel<<<(unsigned) blocks, 256, 0, stream>>>(nnz, prow, pcol, trans, ro, co, K, N_cols, lc_shift, nlc, cnt);
        spd_scan_partial_kernel<<<(unsigned) nb, 256, 0, stream>>>(N, cnt, bsum);
        spd_scan_blocks_kernel<<<1, 1, 0, stream>>>(nb, bsum);
        spd_scan_final_kernel<<<(unsigned) nb, 256, 0, stream>>>(N, cnt, bsum, ptr, cursor);
        spd_scatter_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(nnz, prow, pcol, vals, trans, ro, co, K, N_cols, lc_shift, nlc, cursor, tmp);
        int64_t sb = (N * 32 + 255) / 256;
        if (sb > kNumSMs * 32) sb = kNumSMs * 32;
        if (sb < 1) sb = 1;
        spd_segment_sort_kernel<T><<<(unsigned) sb, 256, 0, stream>>>(N, ptr, tmp, out_l, out_v, K < 0x7fffffffll);
        for (int q = 0; q < 6; ++q) count_launch();
        rc = check_launch("spd_group_by_column");
    }
    if (cnt) cudaFreeAsync(cnt, stream);
    if (cursor) cudaFreeAsync(cursor, stream);
    if (bsum) cudaFreeAsync(bsum, stream);
    if (tmp) cudaFreeAsync(tmp, stream);
    if (rc != RBX_OK) {
        if (ptr) cudaFreeAsync(ptr, stream);
        if (out_l) cudaFreeAsync(out_l, stream);
        if (out_v) cudaFreeAsync(out_v, stream);
        return rc;
    }
    *ptr_out = ptr; *lidx_out = out_l; *vals_out = out_v;
    return RBX_OK;
}

#define RBX_INST_SPD(T)                                                                                                          \
    template int launch_spdata_gather<T>(int64_t, const int64_t*, const int64_t*, const T*, int64_t, int64_t, int64_t, int64_t,   \
                                         const float*, bool, bool, T, T, T*, int64_t, int64_t, cudaStream_t);                    \
    template int spd_group_by_column<T>(int64_t, const int64_t*, const int64_t*, const T*, bool, int64_t, int64_t, int64_t,       \
                                        int64_t, int64_t**, int64_t**, T**, cudaStream_t, int, int64_t);
RBX_INST_SPD(float) RBX_INST_SPD(double)
#undef RBX_INST_SPD

}  // namespace rbx
