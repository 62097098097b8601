// K2 (general path): generate-and-contract for any layout / transposition / offset / size, FP32 or FP64.
//
// Replaces dense::lskge3 / dense::rskge3 (skge.hh:173-215, 320-364): the reference materialises the
// needed submatrix of S (submatrix_as_blackbox, dense_skops.hh:594-602) and calls blas::gemm. Here the
// S tile is produced from the counter inside the kernel (shared memory only) and contracted with a
// register-tiled FMA loop, so nothing proportional to d*m is ever written to HBM.
// This is the shape-agnostic kernel (ragged edges, unaligned offsets, vectors, tiny problems); the
// tensor-core kernels in rbx_sketch_dense_tc.cu take the large aligned cases.
//
// A right sketch B = op(A) * op(S) is evaluated as the left sketch B^T = op(S)^T * op(A)^T by swapping
// strides, so one kernel covers both sides.
#include "rbx_kernels.cuh"

namespace rbx {

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;

template <typename T> __device__ __forceinline__ T op_value(float v, bool uniform);
template <> __device__ __forceinline__ float op_value<float>(float v, bool uniform) {
    return uniform ? __fmul_rn(v, (float) 1.7320508075688772) : v;
}
template <> __device__ __forceinline__ double op_value<double>(float v, bool uniform) {
    return uniform ? __dmul_rn((double) v, 1.7320508075688772) : (double) v;
}

template <typename T, char FAMILY, bool EXPLICIT>
__global__ void __launch_bounds__(256) sketch_dense_simt_kernel(const GemmOperands<T> g) {
    __shared__ double tab[32];
    __shared__ T Ss[BK][BM + 4];
    __shared__ T Xs[BK][BN + 4];
    if (FAMILY == 'G') load_logf_tab(tab);
    const int tid = threadIdx.x;
    const int64_t i0 = ((int64_t) blockIdx.x / g.nbx) * BM, j0 = ((int64_t) blockIdx.x % g.nbx) * BN;
    const int ty = tid / 16, tx = tid % 16;
    T acc[TM][TN];
#pragma unroll
    for (int a = 0; a < TM; ++a)
#pragma unroll
        for (int b = 0; b < TN; ++b) acc[a][b] = (T) 0;
    __syncthreads();

    for (int64_t l0 = 0; l0 < g.K; l0 += BK) {
        // ---- operator tile Ss[l][i] ----
        if (EXPLICIT) {
            for (int e = tid; e < BM * BK; e += 256) {
                int i, l;
                if (g.ss_l == 1) { l = e % BK; i = e / BK; } else { i = e % BM; l = e / BM; }
                const int64_t gi = i0 + i, gl = l0 + l;
                Ss[l][i] = (gi < g.M && gl < g.K) ? g.Sx[gi * g.ss_i + gl * g.ss_l] : (T) 0;
            }
        } else {
            // tile in parent coordinates: rows [R0, R0+nR), cols [C0, C0+nC)
            const int64_t R0 = g.w.pr0 + (g.l_on_cols ? i0 : l0), C0 = g.w.pc0 + (g.l_on_cols ? l0 : i0);
            const int nR = g.l_on_cols ? BM : BK, nC = g.l_on_cols ? BK : BM;
            const int64_t limR = g.l_on_cols ? g.M - i0 : g.K - l0, limC = g.l_on_cols ? g.K - l0 : g.M - i0;
            const int64_t b0 = C0 / 4;
            const int nb = (int)((C0 + nC - 1) / 4 - b0) + 1;
            for (int e = tid; e < nR * nb; e += 256) {
                const int r = e / nb, b = e % nb;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < limR) {
                    const Ctr128 c = ctr_add(g.w.ctr, (uint64_t)((R0 + r) * g.w.W + b0 + b));
                    v = generate4<FAMILY>(philox4x32_10(c, g.w.key), tab);
                }
                const float vv[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int64_t cc = (b0 + b) * 4 + q - C0;
                    if (cc >= 0 && cc < nC) {
                        const T val = (r < limR && cc < limC) ? op_value<T>(vv[q], FAMILY == 'U') : (T) 0;
                        if (g.l_on_cols) Ss[cc][r] = val; else Ss[r][cc] = val;
                    }
                }
            }
        }
        // ---- data tile Xs[l][j] ----
        for (int e = tid; e < BN * BK; e += 256) {
            int l, j;
            if (g.sx_l == 1) { l = e % BK; j = e / BK; } else { j = e % BN; l = e / BN; }
            const int64_t gl = l0 + l, gj = j0 + j;
            Xs[l][j] = (gl < g.K && gj < g.N) ? g.X[gl * g.sx_l + gj * g.sx_j] : (T) 0;
        }
        __syncthreads();
#pragma unroll
        for (int l = 0; l < BK; ++l) {
            T a[TM], b[TN];
#pragma unroll
            for (int q = 0; q < TM; ++q) a[q] = Ss[l][ty * TM + q];
#pragma unroll
            for (int q = 0; q < TN; ++q) b[q] = Xs[l][tx * TN + q];
#pragma unroll
            for (int q = 0; q < TM; ++q)
#pragma unroll
                for (int s = 0; s < TN; ++s) acc[q][s] = fma(a[q], b[s], acc[q][s]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < TM; ++q)
#pragma unroll
        for (int s = 0; s < TN; ++s) {
            const int64_t gi = i0 + ty * TM + q, gj = j0 + tx * TN + s;
            if (gi < g.M && gj < g.N) {
                T* c = g.C + gi * g.sc_i + gj * g.sc_j;
                const T prev = (g.beta == (T) 0) ? (T) 0 : g.beta * (*c);   // beta == 0: B is not read (BLAS convention)
                *c = g.alpha * acc[q][s] + prev;
            }
        }
}

template <typename T>
int launch_sketch_dense_simt(const GemmOperands<T>& g_in, cudaStream_t stream) {
    GemmOperands<T> g = g_in;
    if (g.M <= 0 || g.N <= 0) return RBX_OK;
    g.nbx = (g.N + BN - 1) / BN;
    const int64_t nblocks = g.nbx * ((g.M + BM - 1) / BM);
    if (nblocks > 0x7fffffffll) return RBX_ERR_UNSUPPORTED;
    dim3 grid((unsigned) nblocks);
    if (g.Sx)                    sketch_dense_simt_kernel<T, 'U', true><<<grid, 256, 0, stream>>>(g);
    else if (g.w.family == 'G')  sketch_dense_simt_kernel<T, 'G', false><<<grid, 256, 0, stream>>>(g);
    else                         sketch_dense_simt_kernel<T, 'U', false><<<grid, 256, 0, stream>>>(g);
    count_launch();
    return check_launch("sketch_dense_simt_kernel");
}
template int launch_sketch_dense_simt<float>(const GemmOperands<float>&, cudaStream_t);
template int launch_sketch_dense_simt<double>(const GemmOperands<double>&, cudaStream_t);

}  // namespace rbx
