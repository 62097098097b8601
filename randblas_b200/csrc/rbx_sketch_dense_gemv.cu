// K2 (vector form): y (M) = alpha * Sop (M x K) * x (K) + beta * y for an implicit dense operator -- sketch_vector
// (skve.hh:152-176, which forwards to sketch_general with n = 1; the reference materialises Sop and calls gemm).
//
// With one right-hand side every generated sample is used exactly once, so the kernel is the generator (K1) with a dot
// product attached: bound by the generator's issue rate (uniform) / its FP64 work (Gaussian), nothing is written but
// per-chunk partial sums. One work item = one Philox counter = four consecutive parent-column entries of one parent row.
//   * parent rows run along i (l_on_cols): a warp owns (row i, a chunk of 32 * 4 * U contraction indices), lanes along
//     the counters, warp-shuffle reduction, one partial per (chunk, i);
//   * parent rows run along l: a lane owns four consecutive i and walks a chunk of rows l; one partial per (chunk, i).
// A second kernel adds the partials of each i in chunk order (fixed order => run-to-run identical) and applies alpha, beta.
#include <cstdlib>
#include "rbx_kernels.cuh"

namespace rbx {

constexpr int GV_U = 8;                        // counters per lane and chunk (l_on_cols)
constexpr int GV_LC = 32 * 4 * GV_U;           // contraction indices per chunk (l_on_cols): 1024
constexpr int GV_ROWS = 64;                    // parent rows per chunk (parent rows along l)

template <typename T> __device__ __forceinline__ T gv_value(float v, bool uniform);
template <> __device__ __forceinline__ float gv_value<float>(float v, bool uniform) { return uniform ? __fmul_rn(v, (float) 1.7320508075688772) : v; }
template <> __device__ __forceinline__ double gv_value<double>(float v, bool uniform) { return uniform ? __dmul_rn((double) v, 1.7320508075688772) : (double) v; }

struct GvArgs {
    DenseOpWindow w;
    int64_t M, K;
    int64_t sx;                                // stride of x
    int64_t nchunks;
};

// parent(pr0 + i, pc0 + l): grid = (nchunks, ceil(M / 8)), 8 warps per block, warp = one row i
template <typename T, char FAMILY>
__global__ void __launch_bounds__(256) gemv_rows_kernel(const GvArgs a, const T* __restrict__ x, T* __restrict__ partial) {
    __shared__ double tab[32];
    if (FAMILY == 'G') { load_logf_tab(tab); __syncthreads(); }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t i = (int64_t) blockIdx.y * 8 + warp;
    if (i >= a.M) return;
    const int64_t l_begin = (int64_t) blockIdx.x * GV_LC;                 // first window column of this chunk
    const int64_t cb0 = (a.w.pc0 + l_begin) >> 2;                         // first counter column touched
    const int64_t row_ctr = (a.w.pr0 + i) * a.w.W;
    T acc = (T) 0;
    // interior chunks of an aligned window with a unit-stride, 16-byte aligned x (the usual call): no per-element bounds, x read
    // four elements at a time, the counter a running 64-bit sum (its high half is constant unless the low half wraps here)
    const Ctr128 c0 = ctr_add(a.w.ctr, (uint64_t)(row_ctr + cb0 + lane));
    const uint64_t c0_lo = ((uint64_t) c0.v[1] << 32) | c0.v[0];
    if ((a.w.pc0 & 3) == 0 && a.sx == 1 && l_begin + GV_LC <= a.K && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && c0_lo + 32ull * GV_U >= c0_lo) {
        const T* xp = x + l_begin + 4 * lane;
#pragma unroll 2
        for (int u = 0; u < GV_U; ++u) {
            const uint64_t lo = c0_lo + 32ull * u;
            const float4 v = generate4<FAMILY>(philox4x32_10((uint32_t) lo, (uint32_t)(lo >> 32), c0.v[2], c0.v[3], a.w.key.v[0], a.w.key.v[1]), tab);
            T xv[4];
            if (sizeof(T) == 4) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(xp + 128 * u));
                xv[0] = (T) t.x; xv[1] = (T) t.y; xv[2] = (T) t.z; xv[3] = (T) t.w;
            } else {
                const double2 t0 = __ldg(reinterpret_cast<const double2*>(xp + 128 * u)), t1 = __ldg(reinterpret_cast<const double2*>(xp + 128 * u) + 1);
                xv[0] = (T) t0.x; xv[1] = (T) t0.y; xv[2] = (T) t1.x; xv[3] = (T) t1.y;
            }
            acc = fma(gv_value<T>(v.x, FAMILY == 'U'), xv[0], acc);
            acc = fma(gv_value<T>(v.y, FAMILY == 'U'), xv[1], acc);
            acc = fma(gv_value<T>(v.z, FAMILY == 'U'), xv[2], acc);
            acc = fma(gv_value<T>(v.w, FAMILY == 'U'), xv[3], acc);
        }
    } else
#pragma unroll 2
    for (int u = 0; u <= GV_U; ++u) {                                     // one extra counter when pc0 is not a multiple of 4
        const int64_t cb = cb0 + lane + 32 * u;
        const int64_t l0 = cb * 4 - a.w.pc0;                              // window column of lane 0 of this counter
        if (l0 >= l_begin + GV_LC || l0 >= a.K || l0 + 3 < l_begin) continue;
        const float4 v = generate4<FAMILY>(philox4x32_10(ctr_add(a.w.ctr, (uint64_t)(row_ctr + cb)), a.w.key), tab);
        const float vv[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int64_t l = l0 + q;
            if (l >= l_begin && l < l_begin + GV_LC && l < a.K) acc = fma(gv_value<T>(vv[q], FAMILY == 'U'), x[l * a.sx], acc);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if (lane == 0) partial[(int64_t) blockIdx.x * a.M + i] = acc;
}

// parent(pr0 + l, pc0 + i): grid = (ceil(counter columns / 256), nchunks); a thread owns one counter column (four i)
template <typename T, char FAMILY>
__global__ void __launch_bounds__(256) gemv_cols_kernel(const GvArgs a, const T* __restrict__ x, T* __restrict__ partial) {
    __shared__ double tab[32];
    if (FAMILY == 'G') { load_logf_tab(tab); __syncthreads(); }
    const int64_t cb = (a.w.pc0 >> 2) + (int64_t) blockIdx.x * 256 + threadIdx.x;
    const int64_t i0 = cb * 4 - a.w.pc0;                                  // window row index (of y) of lane 0
    if (i0 >= a.M) return;
    const int64_t l_begin = (int64_t) blockIdx.y * GV_ROWS;
    const int64_t l_end = (l_begin + GV_ROWS < a.K) ? l_begin + GV_ROWS : a.K;
    T acc[4] = { (T) 0, (T) 0, (T) 0, (T) 0 };
    for (int64_t l = l_begin; l < l_end; ++l) {
        const float4 v = generate4<FAMILY>(philox4x32_10(ctr_add(a.w.ctr, (uint64_t)((a.w.pr0 + l) * a.w.W + cb)), a.w.key), tab);
        const T xl = x[l * a.sx];
        acc[0] = fma(gv_value<T>(v.x, FAMILY == 'U'), xl, acc[0]);
        acc[1] = fma(gv_value<T>(v.y, FAMILY == 'U'), xl, acc[1]);
        acc[2] = fma(gv_value<T>(v.z, FAMILY == 'U'), xl, acc[2]);
        acc[3] = fma(gv_value<T>(v.w, FAMILY == 'U'), xl, acc[3]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int64_t i = i0 + q;
        if (i >= 0 && i < a.M) partial[(int64_t) blockIdx.y * a.M + i] = acc[q];
    }
}

template <typename T>
__global__ void __launch_bounds__(256) gemv_reduce_kernel(const T* __restrict__ partial, int64_t nchunks, int64_t M, T alpha, T beta,
                                                         T* __restrict__ y, int64_t sy) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    T s = (T) 0;
    for (int64_t c = 0; c < nchunks; ++c) s += partial[c * M + i];
    T* p = y + i * sy;
    *p = (beta == (T) 0) ? alpha * s : alpha * s + beta * (*p);           // beta == 0: y is not read
}

template <typename T>
int launch_sketch_dense_gemv(const GemmOperands<T>& g, cudaStream_t stream) {
    if (g.N != 1 || g.Sx != nullptr || g.M <= 0 || g.K <= 0) return RBX_ERR_UNSUPPORTED;
    if ((double) g.M * (double) g.K < 32768.0 || getenv("RBX_NO_GEMV")) return RBX_ERR_UNSUPPORTED;     // tiny: the SIMT kernel's one launch wins
    if (g.alpha == (T) 0) return launch_scale_view<T>(g.C, g.M, 1, g.sc_i, g.sc_j, g.beta, stream);
    GvArgs a;
    a.w = g.w; a.M = g.M; a.K = g.K; a.sx = g.sx_l;
    a.nchunks = g.l_on_cols ? (g.K + GV_LC - 1) / GV_LC : (g.K + GV_ROWS - 1) / GV_ROWS;
    // grid.y limits: the row kernel puts row blocks there, the column kernel the chunks
    if (g.l_on_cols ? ((g.M + 7) / 8 > 65535 || a.nchunks > 0x7fffffffll) : (a.nchunks > 65535)) return RBX_ERR_UNSUPPORTED;
    T* partial = nullptr;
    if (ws_alloc(&partial, sizeof(T) * (size_t) a.nchunks * (size_t) g.M, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    const bool G = g.w.family == 'G';
    if (g.l_on_cols) {
        dim3 grid((unsigned) a.nchunks, (unsigned)((g.M + 7) / 8));
        if (G) gemv_rows_kernel<T, 'G'><<<grid, 256, 0, stream>>>(a, g.X, partial);
        else   gemv_rows_kernel<T, 'U'><<<grid, 256, 0, stream>>>(a, g.X, partial);
    } else {
        const int64_t ncb = ((g.w.pc0 + g.M + 3) >> 2) - (g.w.pc0 >> 2);
        dim3 grid((unsigned)((ncb + 255) / 256), (unsigned) a.nchunks);
        if (G) gemv_cols_kernel<T, 'G'><<<grid, 256, 0, stream>>>(a, g.X, partial);
        else   gemv_cols_kernel<T, 'U'><<<grid, 256, 0, stream>>>(a, g.X, partial);
    }
    count_launch();
    int rc = check_launch("gemv kernel");
    if (rc == RBX_OK) {
        gemv_reduce_kernel<T><<<(unsigned)((g.M + 255) / 256), 256, 0, stream>>>(partial, a.nchunks, g.M, g.alpha, g.beta, g.C, g.sc_i);
        count_launch();
        rc = check_launch("gemv_reduce_kernel");
    }
    cudaFreeAsync(partial, stream);
    return rc;
}
template int launch_sketch_dense_gemv<float>(const GemmOperands<float>&, cudaStream_t);
template int launch_sketch_dense_gemv<double>(const GemmOperands<double>&, cudaStream_t);

// ---- allocation-free variant (rbx_sketch_vector_noalloc) -------------------------------------------------------------
// Same two kernels, but the per-chunk partial sums live in a file-scope device array instead of a stream-ordered
// allocation, for callers that may not allocate (CUDA graph capture, serving loops). Every GPU that loads the library holds
// the array (GV_STATIC_BYTES) until exit. All calls share it: calls that overlap on the GPU (different streams) race.
// Nothing in it has to start at zero -- the kernels write every (chunk, i) partial before the reduce kernel reads it -- so
// the previous call's contents are simply overwritten.
constexpr size_t GV_STATIC_BYTES = (size_t) 4 << 20;
static __device__ __align__(256) unsigned char g_gv_scratch[GV_STATIC_BYTES];

// In host code `g_gv_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached.
static void* gv_static_scratch() {
    static void* ptr[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return nullptr; }
    if (!ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_gv_scratch);
        if (e != cudaSuccess) {
            std::fprintf(stderr, "[rbx] cudaGetSymbolAddress(g_gv_scratch): %s\n", cudaGetErrorString(e));
            cudaGetLastError();
            return nullptr;
        }
        ptr[dev] = p;
    }
    return ptr[dev];
}

size_t sketch_vector_static_capacity() { return GV_STATIC_BYTES; }

template <typename T>
int launch_sketch_dense_gemv_static(const GemmOperands<T>& g, cudaStream_t stream) {
    if (g.N != 1 || g.Sx != nullptr || g.M <= 0 || g.K <= 0) return RBX_ERR_UNSUPPORTED;
    if (g.alpha == (T) 0) return launch_scale_view<T>(g.C, g.M, 1, g.sc_i, g.sc_j, g.beta, stream);      // one kernel, no workspace
    GvArgs a;
    a.w = g.w; a.M = g.M; a.K = g.K; a.sx = g.sx_l;
    a.nchunks = g.l_on_cols ? (g.K + GV_LC - 1) / GV_LC : (g.K + GV_ROWS - 1) / GV_ROWS;
    if (a.nchunks > 65535 || (g.M + 7) / 8 > 65535 ||
        (double) a.nchunks * (double) g.M * (double) sizeof(T) > (double) GV_STATIC_BYTES) {
        std::fprintf(stderr, "[rbx] rbx_sketch_vector_noalloc: %lld x %lld needs %lld partial sums of %d bytes; the static scratch holds %lld bytes\n",
                     (long long) g.M, (long long) g.K, (long long)(a.nchunks * g.M), (int) sizeof(T), (long long) GV_STATIC_BYTES);
        return RBX_ERR_TOO_LARGE;
    }
    T* partial = static_cast<T*>(gv_static_scratch());
    if (!partial) return RBX_ERR_CUDA;
    const bool G = g.w.family == 'G';
    if (g.l_on_cols) {
        dim3 grid((unsigned) a.nchunks, (unsigned)((g.M + 7) / 8));
        if (G) gemv_rows_kernel<T, 'G'><<<grid, 256, 0, stream>>>(a, g.X, partial);
        else   gemv_rows_kernel<T, 'U'><<<grid, 256, 0, stream>>>(a, g.X, partial);
    } else {
        const int64_t ncb = ((g.w.pc0 + g.M + 3) >> 2) - (g.w.pc0 >> 2);
        dim3 grid((unsigned)((ncb + 255) / 256), (unsigned) a.nchunks);
        if (G) gemv_cols_kernel<T, 'G'><<<grid, 256, 0, stream>>>(a, g.X, partial);
        else   gemv_cols_kernel<T, 'U'><<<grid, 256, 0, stream>>>(a, g.X, partial);
    }
    count_launch();
    int rc = check_launch("gemv kernel (static scratch)");
    if (rc == RBX_OK) {
        gemv_reduce_kernel<T><<<(unsigned)((g.M + 255) / 256), 256, 0, stream>>>(partial, a.nchunks, g.M, g.alpha, g.beta, g.C, g.sc_i);
        count_launch();
        rc = check_launch("gemv_reduce_kernel (static scratch)");
    }
    return rc;
}
template int launch_sketch_dense_gemv_static<float>(const GemmOperands<float>&, cudaStream_t);
template int launch_sketch_dense_gemv_static<double>(const GemmOperands<double>&, cudaStream_t);

}  // namespace rbx
