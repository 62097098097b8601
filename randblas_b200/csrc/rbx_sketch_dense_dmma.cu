// K2 (double precision, the C5 case): panelled generate-and-contract on the FP64 tensor cores.
//
// Replaces dense::lskge3 / dense::rskge3 for T = double (skge.hh:173-215, 320-364: materialise submat(S), blas::gemm).
// The FP64 pipe is the bound here (DMMA.8x8x4 is the only FP64 tensor instruction sm_100a has: measured 37.1 TFLOP/s
// register-resident, cuBLAS DGEMM 35.5), and the bit-exact Box-Muller also runs on that pipe (~50 FP64 operations per
// sample). Generating the operator tile inside every CTA would repeat that work N / 128 times (39 % extra FP64 work at
// a 128-wide tile), so the operator is generated ONCE per k-panel into a fixed-size, L2-resident workspace
// (M x Kp doubles, <= 256 MB whatever d * m is) by K1 and contracted by the DMMA kernel below, panel after panel with
// beta = 1. Nothing proportional to d * m is allocated; generation is ~1 % of the FP64 work.
//
// dmma_gemm_kernel: C (M x N) = alpha * Sp (M x Kp, k contiguous) * X (Kp x N) + beta * C.
//   CTA tile 128 x 128 x 32, 8 warps as 2 (M) x 4 (N), warp tile 64 x 32 = 8 x 4 DMMA.8x8x4 fragments (64 accumulators
//   per thread). Operands come through a 3-stage cp.async ring (16-byte chunks, zero-fill for ragged edges; 8-byte chunks
//   for an operand whose base or leading dimension is not 16-byte aligned); rows are padded so that the 8-byte fragment
//   loads of a half-warp hit 16 distinct bank pairs.
#include "rbx_kernels.cuh"
#include "rbx_copy.cuh"
#include "rbx_tma.cuh"
#include <cstdlib>

#ifndef RBX_DMMA_MIN_N
#define RBX_DMMA_MIN_N 2         // tall-skinny data: with split-K the padded 128-column tile beats the SIMT kernel
#endif
#ifndef RBX_DMMA_MBAR
#define RBX_DMMA_MBAR 1          // 0: the former pipeline (cp.async groups + one __syncthreads per k-tile)
#endif
#ifndef RBX_DMMA_MIN_K
#define RBX_DMMA_MIN_K 16        // a short contraction axis (lifts): one k-tile is enough to beat the SIMT kernel
#endif
namespace rbx {

#ifndef RBX_DMMA_BK
#define RBX_DMMA_BK 32          // measured at the C5 slab: 16 x 4 stages 33.2, 16 x 5 33.4, 32 x 3 33.6, 32 x 2 28.7 TFLOP/s
#endif
#ifndef RBX_DMMA_STAGES
#define RBX_DMMA_STAGES 3
#endif
constexpr int DM_BM = 128, DM_BN = 128, DM_BK = RBX_DMMA_BK, DM_STAGES = RBX_DMMA_STAGES, DM_THREADS = 256;
constexpr int DM_CH = DM_BK / 2;             // 16-byte chunks per k-contiguous tile row
constexpr int DM_PITCH_K = DM_BK + 4;        // k-contiguous tile: 128 rows x (BK + 4) doubles (row shift 32 B mod 128 B)
constexpr int DM_PITCH_J = DM_BN + 4;        // j-contiguous data tile: 16 rows x 132 doubles
constexpr int DM_TILE_DOUBLES = DM_BM * DM_PITCH_K;                 // 2560 doubles = 20480 B at BK = 16 (>= BK * 132)
static_assert(DM_BM * DM_PITCH_K >= DM_BK * DM_PITCH_J, "the j-contiguous data tile must fit the operand slot");
constexpr size_t DM_SMEM = (size_t) DM_STAGES * 2 * DM_TILE_DOUBLES * sizeof(double);   // 163840 B at BK = 16, 4 stages
static_assert(DM_SMEM >= (size_t) 128 * 136 * sizeof(double) && DM_SMEM <= 227 * 1024, "ring too small for the epilogue staging or too large for an SM");

struct DmmaArgs {
    const double* S; int64_t s_pitch;        // operator panel, S[i * s_pitch + l]
    const double* X; int64_t sx_l, sx_j;     // data, X[l * sx_l + j * sx_j]; one of the strides is 1
    double* C; int64_t sc_i, sc_j;
    int64_t M, N, K;                         // K = panel length
    double alpha, beta;
    int tiles_n;
    // split-K for shapes with few output tiles (tall-skinny data): CTA (tile, ks) contracts k in [ks * kper, (ks + 1) * kper) and
    // writes its tile of partial[ks] (M x N, unit stride along C's unit-stride axis); dmma_reduce_kernel adds them in order
    int ksplit; int64_t kper, ntiles;
    double* partial;
    // 16-byte copies need a 16-byte aligned base and an even pitch; an operand without them (an odd leading dimension) is
    // staged with 8-byte copies instead: twice the copy instructions, the same DMMA stream
    int s16, x16;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src, int src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__device__ __forceinline__ int clamp_bytes(int64_t elems) {
    return elems >= 2 ? 16 : (elems == 1 ? 8 : 0);
}

// XJ = false: X is k-contiguous (sx_l == 1; column-major data, NoTrans). XJ = true: X is j-contiguous (sx_j == 1).
// MF = 8-row fragments per warp along M: the CTA tile is (16 * MF) x 128. The launcher picks MF in {8, 7, 6} (tile heights
// 128 / 112 / 96) so that the number of CTAs fills whole waves of the 148 SMs.
template <bool XJ, bool CI, int MF>
__global__ void __launch_bounds__(DM_THREADS, 1) dmma_gemm_kernel(const DmmaArgs a) {
    constexpr int BM = 16 * MF;
    extern __shared__ __align__(16) double dm_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;                 // 2 x 4 warps
    const int g = lane >> 2, t = lane & 3;
    const int64_t tile_id = (int64_t) blockIdx.x % a.ntiles, ks = (int64_t) blockIdx.x / a.ntiles;
    const int64_t i0 = (tile_id / a.tiles_n) * BM, j0 = (tile_id % a.tiles_n) * DM_BN;
    const uint32_t smem_base = (uint32_t) __cvta_generic_to_shared(dm_smem);
    // this CTA's slice of the contraction axis (the whole panel when ksplit == 1); kper is a multiple of 16, so the slice
    // keeps the 16-byte alignment of both operands
    const int64_t kbeg = ks * a.kper;
    const int64_t Kc = (a.ksplit > 1) ? ((kbeg + a.kper < a.K) ? a.kper : a.K - kbeg) : a.K;
    const double* __restrict__ Sb = a.S + kbeg;
    const double* __restrict__ Xb = a.X + kbeg * a.sx_l;
    const int nk = (int)((Kc + DM_BK - 1) / DM_BK);

    auto load_stage = [&](int kt, int slot) {
        const int64_t k0 = (int64_t) kt * DM_BK;
        const uint32_t sS = smem_base + (uint32_t)(slot * 2 * DM_TILE_DOUBLES) * 8u;
        const uint32_t sX = sS + (uint32_t) DM_TILE_DOUBLES * 8u;
        if (a.s16) {
#pragma unroll
            for (int q = 0; q < (BM * DM_CH + DM_THREADS - 1) / DM_THREADS; ++q) {
                const int c = tid + q * DM_THREADS, row = c / DM_CH, kc = (c % DM_CH) * 2;
                if (BM * DM_CH % DM_THREADS != 0 && row >= BM) break;
                const int64_t gi = i0 + row;
                const bool ok = gi < a.M;
                const double* src = Sb + (ok ? gi : 0) * a.s_pitch + (k0 + kc < Kc ? k0 + kc : 0);
                cp_async16(sS + (uint32_t)(row * DM_PITCH_K + kc) * 8u, src, ok ? clamp_bytes(Kc - k0 - kc) : 0);
            }
        } else {
#pragma unroll
            for (int q = 0; q < BM * DM_BK / DM_THREADS; ++q) {
                const int c = tid + q * DM_THREADS, row = c / DM_BK, kc = c % DM_BK;
                const bool ok = i0 + row < a.M && k0 + kc < Kc;
                cp_async8(sS + (uint32_t)(row * DM_PITCH_K + kc) * 8u, ok ? Sb + (i0 + row) * a.s_pitch + k0 + kc : Sb, ok ? 8 : 0);
            }
        }
        if (!XJ) {
            if (a.x16) {
#pragma unroll
                for (int q = 0; q < DM_BN * DM_CH / DM_THREADS; ++q) {
                    const int c = tid + q * DM_THREADS, row = c / DM_CH, kc = (c % DM_CH) * 2;
                    const int64_t gj = j0 + row;
                    const bool ok = gj < a.N;
                    const double* src = Xb + (ok ? gj : 0) * a.sx_j + (k0 + kc < Kc ? k0 + kc : 0);
                    cp_async16(sX + (uint32_t)(row * DM_PITCH_K + kc) * 8u, src, ok ? clamp_bytes(Kc - k0 - kc) : 0);
                }
            } else {
#pragma unroll
                for (int q = 0; q < DM_BN * DM_BK / DM_THREADS; ++q) {
                    const int c = tid + q * DM_THREADS, row = c / DM_BK, kc = c % DM_BK;
                    const bool ok = j0 + row < a.N && k0 + kc < Kc;
                    cp_async8(sX + (uint32_t)(row * DM_PITCH_K + kc) * 8u, ok ? Xb + (j0 + row) * a.sx_j + k0 + kc : Xb, ok ? 8 : 0);
                }
            }
        } else {
            if (a.x16) {
#pragma unroll
                for (int q = 0; q < DM_BK * 64 / DM_THREADS; ++q) {
                    const int c = tid + q * DM_THREADS, l = c >> 6, jc = (c & 63) * 2;
                    const int64_t gl = k0 + l;
                    const bool ok = gl < Kc;
                    const double* src = Xb + (ok ? gl : 0) * a.sx_l + (j0 + jc < a.N ? j0 + jc : 0);
                    cp_async16(sX + (uint32_t)(l * DM_PITCH_J + jc) * 8u, src, ok ? clamp_bytes(a.N - j0 - jc) : 0);
                }
            } else {
#pragma unroll
                for (int q = 0; q < DM_BK * 128 / DM_THREADS; ++q) {
                    const int c = tid + q * DM_THREADS, l = c >> 7, jc = c & 127;
                    const bool ok = k0 + l < Kc && j0 + jc < a.N;
                    cp_async8(sX + (uint32_t)(l * DM_PITCH_J + jc) * 8u, ok ? Xb + (k0 + l) * a.sx_l + j0 + jc : Xb, ok ? 8 : 0);
                }
            }
        }
    };

    double acc[MF][4][2];
#pragma unroll
    for (int f = 0; f < MF; ++f)
#pragma unroll
        for (int h = 0; h < 4; ++h) { acc[f][h][0] = 0.0; acc[f][h][1] = 0.0; }

#if RBX_DMMA_MBAR
    // The ring is handed over with two mbarriers per stage instead of one CTA-wide barrier per k-tile: `full` completes when
    // the cp.async copies of all 256 threads have landed, `empty` when all 8 warps have read the stage. A warp refills the
    // stage of tile kt - 1 AFTER computing tile kt, so it waits for the other warps only if they are a whole tile behind, and
    // the two warps of a scheduler drift apart instead of stalling at the same instruction (the barrier version left the
    // FP64 tensor pipe idle 17 % of the time: barrier + the LDS latency right after it, ncu).
    __shared__ __align__(8) uint64_t full_bar[DM_STAGES], empty_bar[DM_STAGES];
    if (tid == 0) {
        for (int s = 0; s < DM_STAGES; ++s) { mbar_init(&full_bar[s], DM_THREADS); mbar_init(&empty_bar[s], DM_THREADS / 32); }
        fence_mbar_init();
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < DM_STAGES - 1; ++s)
        if (s < nk) {
            load_stage(s, s);
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(smem_u32(&full_bar[s])) : "memory");
        }
    for (int kt = 0; kt < nk; ++kt) {
        mbar_wait(&full_bar[kt % DM_STAGES], (uint32_t)(kt / DM_STAGES) & 1u);
#else
#pragma unroll
    for (int s = 0; s < DM_STAGES - 1; ++s) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nk; ++kt) {
        cp_async_wait<DM_STAGES - 2>();
        __syncthreads();
        if (kt + DM_STAGES - 1 < nk) load_stage(kt + DM_STAGES - 1, (kt + DM_STAGES - 1) % DM_STAGES);
        cp_async_commit();
#endif
        const double* Ss = dm_smem + (size_t)(kt % DM_STAGES) * 2 * DM_TILE_DOUBLES;
        const double* Xs = Ss + DM_TILE_DOUBLES;
        const double* pa = Ss + (wm * (8 * MF) + g) * DM_PITCH_K + t;
        const double* pb = XJ ? Xs + t * DM_PITCH_J + wn * 32 + g : Xs + (wn * 32 + g) * DM_PITCH_K + t;
#pragma unroll
        for (int kk = 0; kk < DM_BK / 4; ++kk) {
            double fa[MF], fb[4];
#pragma unroll
            for (int f = 0; f < MF; ++f) fa[f] = pa[f * 8 * DM_PITCH_K + kk * 4];
#pragma unroll
            for (int h = 0; h < 4; ++h) fb[h] = XJ ? pb[kk * 4 * DM_PITCH_J + h * 8] : pb[h * 8 * DM_PITCH_K + kk * 4];
#pragma unroll
            for (int f = 0; f < MF; ++f)
#pragma unroll
                for (int h = 0; h < 4; ++h) dmma884(acc[f][h][0], acc[f][h][1], fa[f], fb[h]);
        }
#if RBX_DMMA_MBAR
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[kt % DM_STAGES]);
        const int nxt = kt + DM_STAGES - 1;
        if (nxt < nk) {
            const int ls = nxt % DM_STAGES;
            if (nxt >= DM_STAGES) mbar_wait(&empty_bar[ls], (uint32_t)(nxt / DM_STAGES - 1) & 1u);
            load_stage(nxt, ls);
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(smem_u32(&full_bar[ls])) : "memory");
        }
#endif
    }
    cp_async_wait<0>();

    // epilogue: lane (g, t) holds C[row g][cols 2t, 2t + 1] of each 8 x 8 fragment. The tile goes through shared memory
    // (the ring is free now) so that global traffic runs along C's unit-stride axis in 256-byte warp rows, with the
    // old values of four rows in flight per thread when beta != 0.
    __syncthreads();
    constexpr int P = CI ? 130 : 136;            // staging pitch: conflict-free for the fragment stores below
    if (CI) {                                     // C is i-contiguous (ColMajor B): stage as Cs[j][i]
#pragma unroll
        for (int f = 0; f < MF; ++f)
#pragma unroll
            for (int h = 0; h < 4; ++h)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    dm_smem[(wn * 32 + h * 8 + 2 * t + e) * P + wm * (8 * MF) + f * 8 + g] = acc[f][h][e];
    } else {                                      // C is j-contiguous: stage as Cs[i][j], two columns per 16-byte store
#pragma unroll
        for (int f = 0; f < MF; ++f)
#pragma unroll
            for (int h = 0; h < 4; ++h)
                *reinterpret_cast<double2*>(&dm_smem[(wm * (8 * MF) + f * 8 + g) * P + wn * 32 + h * 8 + 2 * t]) =
                    make_double2(acc[f][h][0], acc[f][h][1]);
    }
    __syncthreads();
    // "line" = index along the slow axis of C, "pos" = index along its unit-stride axis
    const int64_t line0 = CI ? j0 : i0, pos0 = CI ? i0 : j0;
    const int64_t nline = CI ? a.N : a.M, npos = CI ? a.M : a.N;
    const bool split = a.ksplit > 1;
    double* __restrict__ Cb = split ? a.partial + (size_t) ks * (size_t) a.M * (size_t) a.N : a.C;
    const int64_t s_line = split ? npos : (CI ? a.sc_j : a.sc_i);
    const double alpha = split ? 1.0 : a.alpha, beta = split ? 0.0 : a.beta;
#pragma unroll 1
    for (int lb = warp * 16; lb < warp * 16 + 16; lb += 4) {
        double old[4][4];
        if (beta != 0.0) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int64_t gl = line0 + lb + u, gp = pos0 + lane + 32 * q;
                    const bool in_tile = CI ? (lane + 32 * q < BM) : (lb + u < BM);
                    old[u][q] = (in_tile && gl < nline && gp < npos) ? Cb[gl * s_line + gp] : 0.0;
                }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t gl = line0 + lb + u, gp = pos0 + lane + 32 * q;
                const bool in_tile = CI ? (lane + 32 * q < BM) : (lb + u < BM);
                if (in_tile && gl < nline && gp < npos) {
                    const double v = alpha * dm_smem[(lb + u) * P + lane + 32 * q];
                    Cb[gl * s_line + gp] = (beta == 0.0) ? v : v + beta * old[u][q];   // beta == 0: C is not read
                }
            }
    }
}

// C = alpha * sum over ks of partial[ks] + beta * C, the splits added in increasing order (fixed => run-to-run identical);
// partial[ks] is M x N with unit stride along C's unit-stride axis (ci: along i)
__global__ void __launch_bounds__(256) dmma_reduce_kernel(const double* __restrict__ partial, int ksplit, int64_t M, int64_t N, bool ci,
                                                        double alpha, double beta, double* __restrict__ C, int64_t sc_i, int64_t sc_j) {
    const int64_t total = M * N;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        const int64_t line = ci ? e / M : e / N, pos = ci ? e - line * M : e - line * N;
        const int64_t i = ci ? pos : line, j = ci ? line : pos;
        double sum = 0.0;
        for (int p = 0; p < ksplit; ++p) sum += partial[(size_t) p * total + e];
        double* c = C + i * sc_i + j * sc_j;
        *c = (beta == 0.0) ? alpha * sum : alpha * sum + beta * (*c);
    }
}

template <int MF>
static int launch_dmma_gemm_mf(const DmmaArgs& a, int64_t tiles, cudaStream_t stream) {
    const bool xj = (a.sx_l != 1), ci = (a.sc_i == 1);
    auto kern = xj ? (ci ? dmma_gemm_kernel<true, true, MF> : dmma_gemm_kernel<true, false, MF>)
                   : (ci ? dmma_gemm_kernel<false, true, MF> : dmma_gemm_kernel<false, false, MF>);
    static bool attr_set[4] = { false, false, false, false };
    const int which = (xj ? 2 : 0) + (ci ? 1 : 0);
    if (!attr_set[which]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) DM_SMEM) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_CUDA; }
        attr_set[which] = true;
    }
    kern<<<(unsigned)(tiles * a.ksplit), DM_THREADS, DM_SMEM, stream>>>(a);
    count_launch();
    return check_launch("dmma_gemm_kernel");
}

// tile height for this shape: the one whose CTA count wastes the least of its last wave (cost = waves x tile height)
static int pick_mf(int64_t M, int64_t N) {
    if (const char* ev = getenv("RBX_DMMA_MF")) { const int v = atoi(ev); if (v >= 6 && v <= 8) return v; }
    const int64_t tn = (N + DM_BN - 1) / DM_BN;
    int best = 8; int64_t best_cost = -1;
    for (int mf = 8; mf >= 6; --mf) {
        const int64_t tiles = tn * ((M + 16 * mf - 1) / (16 * mf));
        const int64_t cost = ((tiles + kNumSMs - 1) / kNumSMs) * mf;
        if (best_cost < 0 || cost < best_cost) { best = mf; best_cost = cost; }
    }
    return best;
}

static int launch_dmma_gemm(const DmmaArgs& a_in, int mf, cudaStream_t stream) {
    DmmaArgs a = a_in;
    a.tiles_n = (int)((a.N + DM_BN - 1) / DM_BN);
    const int64_t tiles = (int64_t) a.tiles_n * ((a.M + 16 * mf - 1) / (16 * mf));
    if (tiles > 0x7fffffffll) return RBX_ERR_UNSUPPORTED;
    // few output tiles (tall-skinny data: M x N small, K long): split the contraction axis so that about two waves of CTAs run
    int64_t ksplit = (2 * kNumSMs) / tiles;
    const int64_t nk = (a.K + DM_BK - 1) / DM_BK;
    if (ksplit > nk / (256 / DM_BK)) ksplit = nk / (256 / DM_BK);   // at least 256 contraction indices per split
    if (ksplit < 1) ksplit = 1;
    if (const char* ev = getenv("RBX_DMMA_KSPLIT")) { const int64_t v = atoll(ev); if (v >= 1 && v <= nk) ksplit = v; }
    a.ntiles = tiles; a.ksplit = (int) ksplit; a.partial = nullptr;
    a.s16 = reinterpret_cast<uintptr_t>(a.S) % 16 == 0 && a.s_pitch % 2 == 0;
    a.x16 = reinterpret_cast<uintptr_t>(a.X) % 16 == 0 && (a.sx_l == 1 ? a.sx_j : a.sx_l) % 2 == 0;
    a.kper = ((nk + ksplit - 1) / ksplit) * DM_BK;
    a.ksplit = (int)((a.K + a.kper - 1) / a.kper);            // no empty splits
    if (a.ksplit > 1 && ws_alloc(&a.partial, sizeof(double) * (size_t) a.ksplit * (size_t) a.M * (size_t) a.N, stream) != cudaSuccess) {
        cudaGetLastError();
        return RBX_ERR_ALLOC;
    }
    int rc = mf == 8 ? launch_dmma_gemm_mf<8>(a, tiles, stream) : mf == 7 ? launch_dmma_gemm_mf<7>(a, tiles, stream)
                                                                        : launch_dmma_gemm_mf<6>(a, tiles, stream);
    if (rc == RBX_OK && a.ksplit > 1) {
        int64_t blocks = (a.M * a.N + 255) / 256;
        if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
        dmma_reduce_kernel<<<(unsigned) blocks, 256, 0, stream>>>(a.partial, a.ksplit, a.M, a.N, a.sc_i == 1, a.alpha, a.beta, a.C, a.sc_i, a.sc_j);
        count_launch();
        rc = check_launch("dmma_reduce_kernel");
    }
    if (a.partial) cudaFreeAsync(a.partial, stream);
    return rc;
}

int launch_sketch_dense_tc_f64(const GemmOperands<double>& g, cudaStream_t stream) {
    if (getenv("RBX_NO_TC")) return RBX_ERR_UNSUPPORTED;
    if (g.M < 1 || g.N < RBX_DMMA_MIN_N || g.K < RBX_DMMA_MIN_K || (double) g.M * g.N * g.K < dense_min_macs(1.3e8)) return RBX_ERR_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(g.X) % 8 != 0 || (g.sx_l != 1 && g.sx_j != 1)) return RBX_ERR_UNSUPPORTED;
    if (g.Sx && ((g.ss_l != 1 && g.ss_i != 1) || reinterpret_cast<uintptr_t>(g.Sx) % 8 != 0)) return RBX_ERR_UNSUPPORTED;
    if (g.sc_i != 1 && g.sc_j != 1) return RBX_ERR_UNSUPPORTED;
    if (g.alpha == 0.0) return launch_scale_view<double>(g.C, g.M, g.N, g.sc_i, g.sc_j, g.beta, stream);

    DmmaArgs a;
    a.X = g.X; a.sx_l = g.sx_l; a.sx_j = g.sx_j;
    a.C = g.C; a.sc_i = g.sc_i; a.sc_j = g.sc_j;
    a.M = g.M; a.N = g.N; a.alpha = g.alpha; a.tiles_n = 0;
    if (g.Sx && g.ss_l == 1) {                    // DenseSkOp::buff given, k contiguous: one pass over the caller's array
        a.S = g.Sx; a.s_pitch = g.ss_i; a.K = g.K; a.beta = g.beta;
        return launch_dmma_gemm(a, pick_mf(g.M, g.N), stream);
    }
    if (g.Sx) {                                   // ... with unit stride along i (a short-axis-major operator): transposed panel by
        int64_t kp = (((int64_t) 256 << 20) / (8 * g.M)) / 16 * 16;      // panel into a k-contiguous workspace (<= 256 MB) first
        if (kp < 64) kp = 64;
        if (kp > g.K) kp = (g.K + 15) / 16 * 16;
        double* ws = nullptr;
        if (ws_alloc(&ws, sizeof(double) * (size_t) g.M * (size_t) kp, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
        const int mf = pick_mf(g.M, g.N);
        int rc = RBX_OK;
        for (int64_t l0 = 0; l0 < g.K && rc == RBX_OK; l0 += kp) {
            const int64_t kl = (l0 + kp <= g.K) ? kp : g.K - l0;
            rc = launch_kmajor_copy<double>(g.Sx + l0 * g.ss_l, kl, g.M, g.ss_l, g.ss_i, ws, kp, stream);
            if (rc != RBX_OK) break;
            a.S = ws; a.s_pitch = kp; a.K = kl; a.X = g.X + l0 * g.sx_l; a.beta = (l0 == 0) ? g.beta : 1.0;
            rc = launch_dmma_gemm(a, mf, stream);
        }
        cudaFreeAsync(ws, stream);
        return rc;
    }
    // implicit operator: k-panels of a fixed-size workspace (<= 256 MB; the CTAs of one row of tiles run together and share
    // their operator rows through L2). Longer panels amortise the read-modify-write of C between panels.
    int64_t kp = ((int64_t) 256 << 20) / (8 * g.M);
    if (const char* ev = getenv("RBX_DMMA_PANEL")) { const int64_t v = atoll(ev); if (v >= 16) kp = v; }
    kp = (kp / 16) * 16;
    if (kp < 64) kp = 64;
    if (kp > g.K) kp = (g.K + 15) / 16 * 16;
    const int mf = pick_mf(g.M, g.N);
    // Two panel buffers when there is more than one panel: panel p + 1 is generated on an auxiliary stream while the GEMM
    // of panel p runs (the generator is issue-bound, the GEMM DMMA-bound; without the overlap the fill is ~9 % of the time
    // of a 512-column shard). Events order buffer reuse; the caller's stream waits for every fill, so freeing on it is safe.
    const int64_t npanels = (g.K + kp - 1) / kp;
    const int nbuf = (npanels > 1 && !getenv("RBX_DMMA_NO_OVERLAP")) ? 2 : 1;
    const size_t panel_elems = (size_t) g.M * (size_t) kp;
    double* panel = nullptr;
    if (ws_alloc(&panel, sizeof(double) * panel_elems * nbuf, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    cudaStream_t aux = stream;
    cudaEvent_t ready = nullptr, filled[2] = { nullptr, nullptr }, used[2] = { nullptr, nullptr };
    int rc = RBX_OK;
    if (nbuf == 2) {
        aux = aux_stream();
        bool ok = aux != nullptr && cudaEventCreateWithFlags(&ready, cudaEventDisableTiming) == cudaSuccess;
        for (int b = 0; b < 2 && ok; ++b)
            ok = cudaEventCreateWithFlags(&filled[b], cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&used[b], cudaEventDisableTiming) == cudaSuccess;
        ok = ok && cudaEventRecord(ready, stream) == cudaSuccess && cudaStreamWaitEvent(aux, ready, 0) == cudaSuccess;
        if (!ok) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    }
    int64_t p = 0;
    for (int64_t l0 = 0; l0 < g.K && rc == RBX_OK; l0 += kp, ++p) {
        const int64_t kl = (l0 + kp <= g.K) ? kp : g.K - l0;
        const int b = (nbuf == 2) ? (int)(p & 1) : 0;
        double* buf = panel + (size_t) b * panel_elems;
        if (nbuf == 2 && p >= 2 && cudaStreamWaitEvent(aux, used[b], 0) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; break; }
        if (g.l_on_cols)   // parent rows = i, parent columns = l: counters run along the panel's contiguous axis
            rc = launch_fill_window<double>(g.w.family, g.w.W * 4, buf, g.M, kl, g.w.pr0, g.w.pc0 + l0, g.w.ctr, g.w.key,
                                            kp, 1, g.w.family == 'U', aux);
        else               // parent rows = l, parent columns = i
            rc = launch_fill_window<double>(g.w.family, g.w.W * 4, buf, kl, g.M, g.w.pr0 + l0, g.w.pc0, g.w.ctr, g.w.key,
                                            1, kp, g.w.family == 'U', aux);
        if (rc != RBX_OK) break;
        if (nbuf == 2 && (cudaEventRecord(filled[b], aux) != cudaSuccess || cudaStreamWaitEvent(stream, filled[b], 0) != cudaSuccess)) {
            cudaGetLastError(); rc = RBX_ERR_CUDA; break;
        }
        a.S = buf; a.s_pitch = kp; a.K = kl;
        a.X = g.X + l0 * g.sx_l;
        a.beta = (l0 == 0) ? g.beta : 1.0;
        rc = launch_dmma_gemm(a, mf, stream);
        if (rc == RBX_OK && nbuf == 2 && cudaEventRecord(used[b], stream) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    }
    if (nbuf == 2) {
        if (rc != RBX_OK) cudaStreamSynchronize(aux);      // error path: nothing may still be writing the panels
        if (ready) cudaEventDestroy(ready);
        for (int b = 0; b < 2; ++b) { if (filled[b]) cudaEventDestroy(filled[b]); if (used[b]) cudaEventDestroy(used[b]); }
    }
    cudaFreeAsync(panel, stream);
    return rc;
}

}  // namespace rbx
