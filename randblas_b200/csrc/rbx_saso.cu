// K4: streaming SASO sketch  B (d x n) = alpha * S (d x m) * A (m x n) + beta * B,  S = wide Short-axis sparse operator
// with vec_nnz signed ones per column, rows of A contiguous (RowMajor A, or ColMajor A^T).
//
// Replaces sparse::lskges -> left_spmm -> coo::apply_coo_left_jki_p11 (skge.hh:485-510, spmm_dispatch.hh:48-160,
// coo_spmm_impl.hh:79-162): the reference sorts the triplets on the host and makes n passes over them. Here A is read
// from HBM exactly once and B is written once (the algorithmic bytes of SURVEY 8d):
//   * the output tile of a CTA (512 rows x 32 columns) lives in REGISTERS for the whole pass: warp w owns 32 rows,
//     lane l owns column l, so an update is one shared-memory load + one DADD and no atomics;
//   * A streams through a 4-stage shared-memory ring filled by TMA (cp.async.bulk.tensor); the CTAs of a cluster own
//     different row blocks of the same 32 columns and share every A chunk through TMA multicast, so L2/HBM see each
//     byte of A once;
//   * the operator is regenerated on the device (bit-exact Fisher-Yates, rbx_fy.cuh) by a small prep kernel that
//     buckets each 128-column chunk of S by owning warp, in a deterministic order => run-to-run identical results.
// Bound: HBM for float; for double the shared-memory gather (8 B per update, 128 B/clk/SM) is the tighter bound.
#include <cstdlib>
#include "rbx_kernels.cuh"
#include "rbx_fy.cuh"
#include "rbx_tma.cuh"

namespace rbx {

constexpr int SASO_MC = 128;        // rows of A per chunk (= columns of S per chunk)
constexpr int SASO_W = 32;          // columns of A per CTA (one per lane)
constexpr int SASO_WARPS = 16;      // consumer warps per CTA
constexpr int SASO_RW = 32;         // rows of B per consumer warp
constexpr int SASO_NST = 4;         // ring stages
constexpr int SASO_THREADS = SASO_WARPS * 32;   // 512 threads => 128 registers per thread

// ---- prep: bucket the entries of each chunk by owning warp ------------------------------------------------
// ent[chunk * MC * k + off[chunk][b] ...] = 16-bit codes (c_local | r_local << 7 | negative << 12 | valid << 13), in (column, draw) order
template <int KMAX>
__global__ void __launch_bounds__(SASO_MC) saso_prep_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t ro, int64_t co,
                                                          int64_t M, int64_t K, int nbuckets, int off_stride, uint16_t* __restrict__ ent,
                                                          uint32_t* __restrict__ off) {
    extern __shared__ uint16_t sh[];
    uint16_t* code = sh;                               // MC * vec_nnz
    uint16_t* bkt = sh + SASO_MC * vec_nnz;            // MC * vec_nnz
    __shared__ uint32_t cnt[129];
    const int t = threadIdx.x;
    const int64_t chunk = blockIdx.x;
    const int64_t l = chunk * SASO_MC + t;
    for (int j = 0; j < vec_nnz; ++j) bkt[t * vec_nnz + j] = 0xFFFF;
    if (l < K) {
        fy_vector<KMAX>(ctr, key, co + l, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
            const int64_t r = idx - ro;
            if (r >= 0 && r < M) {
                code[t * vec_nnz + j] = (uint16_t)(t | ((int)(r % SASO_RW) << 7) | (negative ? 1 << 12 : 0) | 0x2000);
                bkt[t * vec_nnz + j] = (uint16_t)(r / SASO_RW);
            }
        });
    }
    __syncthreads();
    const int total = SASO_MC * vec_nnz;
    uint32_t mine = 0;
    if (t < nbuckets)
        for (int e = 0; e < total; ++e) mine += (bkt[e] == t);
    if (t < nbuckets) cnt[t + 1] = mine;
    if (t == 0) cnt[0] = 0;
    __syncthreads();
    if (t == 0) for (int b = 0; b < nbuckets; ++b) cnt[b + 1] += cnt[b];
    __syncthreads();
    uint32_t* o = off + chunk * off_stride;
    for (int b = t; b <= nbuckets; b += SASO_MC) o[b] = cnt[b];
    if (t < nbuckets) {
        uint16_t* dst = ent + chunk * total + cnt[t];
        for (int e = 0; e < total; ++e)
            if (bkt[e] == t) *dst++ = code[e];
    }
}

template <typename T>
struct SasoArgs {
    const uint16_t* ent; const uint32_t* off;
    int vec_nnz, nbuckets, njb, msplit, nchunks, nrb, dbg_skip;
    int64_t M, N;
    T alpha, beta;
    T* C; int64_t sc_i, sc_j;
    T* partial;          // [msplit][M][njb * 32] when msplit > 1
};

template <typename T> __device__ __forceinline__ T flip_sign(T v, uint32_t neg);
template <> __device__ __forceinline__ double flip_sign<double>(double v, uint32_t neg) {
    return __hiloint2double(__double2hiint(v) ^ (int)(neg << 31), __double2loint(v));
}
template <> __device__ __forceinline__ float flip_sign<float>(float v, uint32_t neg) {
    return __int_as_float(__float_as_int(v) ^ (int)(neg << 31));
}

// acc[idx] += v for a warp-uniform idx in [0,16): an indirect branch into 16 single-add targets (brx.idx), so the
// accumulators stay in registers. A C++ switch compiles to a 5-level compare/branch tree here, which is issue-bound.
__device__ __forceinline__ void acc_select16(double (&a)[16], uint32_t idx, double v) {
    asm volatile(
        "{\n\t"
        "tl: .branchtargets L0,L1,L2,L3,L4,L5,L6,L7,L8,L9,L10,L11,L12,L13,L14,L15;\n\t"
        "brx.idx.uni %16, tl;\n\t"
        "L0: add.f64 %0, %0, %17; bra.uni LD;\n\t"   "L1: add.f64 %1, %1, %17; bra.uni LD;\n\t"
        "L2: add.f64 %2, %2, %17; bra.uni LD;\n\t"   "L3: add.f64 %3, %3, %17; bra.uni LD;\n\t"
        "L4: add.f64 %4, %4, %17; bra.uni LD;\n\t"   "L5: add.f64 %5, %5, %17; bra.uni LD;\n\t"
        "L6: add.f64 %6, %6, %17; bra.uni LD;\n\t"   "L7: add.f64 %7, %7, %17; bra.uni LD;\n\t"
        "L8: add.f64 %8, %8, %17; bra.uni LD;\n\t"   "L9: add.f64 %9, %9, %17; bra.uni LD;\n\t"
        "L10: add.f64 %10, %10, %17; bra.uni LD;\n\t" "L11: add.f64 %11, %11, %17; bra.uni LD;\n\t"
        "L12: add.f64 %12, %12, %17; bra.uni LD;\n\t" "L13: add.f64 %13, %13, %17; bra.uni LD;\n\t"
        "L14: add.f64 %14, %14, %17; bra.uni LD;\n\t" "L15: add.f64 %15, %15, %17;\n\t"
        "LD:\n\t}\n"
        : "+d"(a[0]), "+d"(a[1]), "+d"(a[2]), "+d"(a[3]), "+d"(a[4]), "+d"(a[5]), "+d"(a[6]), "+d"(a[7]),
          "+d"(a[8]), "+d"(a[9]), "+d"(a[10]), "+d"(a[11]), "+d"(a[12]), "+d"(a[13]), "+d"(a[14]), "+d"(a[15])
        : "r"(idx), "d"(v));
}
__device__ __forceinline__ void acc_select16(float (&a)[16], uint32_t idx, float v) {
    asm volatile(
        "{\n\t"
        "tl: .branchtargets L0,L1,L2,L3,L4,L5,L6,L7,L8,L9,L10,L11,L12,L13,L14,L15;\n\t"
        "brx.idx.uni %16, tl;\n\t"
        "L0: add.f32 %0, %0, %17; bra.uni LD;\n\t"   "L1: add.f32 %1, %1, %17; bra.uni LD;\n\t"
        "L2: add.f32 %2, %2, %17; bra.uni LD;\n\t"   "L3: add.f32 %3, %3, %17; bra.uni LD;\n\t"
        "L4: add.f32 %4, %4, %17; bra.uni LD;\n\t"   "L5: add.f32 %5, %5, %17; bra.uni LD;\n\t"
        "L6: add.f32 %6, %6, %17; bra.uni LD;\n\t"   "L7: add.f32 %7, %7, %17; bra.uni LD;\n\t"
        "L8: add.f32 %8, %8, %17; bra.uni LD;\n\t"   "L9: add.f32 %9, %9, %17; bra.uni LD;\n\t"
        "L10: add.f32 %10, %10, %17; bra.uni LD;\n\t" "L11: add.f32 %11, %11, %17; bra.uni LD;\n\t"
        "L12: add.f32 %12, %12, %17; bra.uni LD;\n\t" "L13: add.f32 %13, %13, %17; bra.uni LD;\n\t"
        "L14: add.f32 %14, %14, %17; bra.uni LD;\n\t" "L15: add.f32 %15, %15, %17;\n\t"
        "LD:\n\t}\n"
        : "+f"(a[0]), "+f"(a[1]), "+f"(a[2]), "+f"(a[3]), "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7]),
          "+f"(a[8]), "+f"(a[9]), "+f"(a[10]), "+f"(a[11]), "+f"(a[12]), "+f"(a[13]), "+f"(a[14]), "+f"(a[15])
        : "r"(idx), "f"(v));
}

// 16-byte asynchronous global->shared copy; bytes beyond src_bytes are zero-filled (ragged edges)
__device__ __forceinline__ void cp_async16(void* dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

template <typename T>
struct SasoStream {
    const T* X; int64_t ldx; int64_t K, N;      // data: K x N, rows contiguous
    int ent_stride;                             // codes per chunk (= MC * vec_nnz, a multiple of 8)
    int off_stride;                             // uint32 per chunk row of `off` (multiple of 4)
};

template <typename T>
__global__ void __launch_bounds__(SASO_THREADS, 1) saso_rows_kernel(const SasoStream<T> x, const SasoArgs<T> a) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int A_BYTES = SASO_MC * SASO_W * (int) sizeof(T);
    const int ent_bytes = x.ent_stride * 2, off_bytes = x.off_stride * 4;
    const int stage_bytes = A_BYTES + ent_bytes + off_bytes;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // the row blocks of one (column block, split) are adjacent block ids: they run together and share A through L2
    const int rb = blockIdx.x % a.nrb;
    const int cid = blockIdx.x / a.nrb;
    const int jb = cid % a.njb;
    const int ms = cid / a.njb;
    const int per = (a.nchunks + a.msplit - 1) / a.msplit;
    const int cb = ms * per;
    const int ce = (cb + per < a.nchunks) ? cb + per : a.nchunks;
    const int64_t j0 = (int64_t) jb * SASO_W;

    auto load_stage = [&](int ch, int s) {
        uint8_t* st = smem + (size_t) s * stage_bytes;
        constexpr int PPR = SASO_W * (int) sizeof(T) / 16;          // 16-byte pieces per row
        constexpr int EP = 16 / (int) sizeof(T);                    // elements per piece
        for (int p = threadIdx.x; p < SASO_MC * PPR; p += SASO_THREADS) {
            const int row = p / PPR, q = p % PPR;
            const int64_t l = (int64_t) ch * SASO_MC + row, col = j0 + (int64_t) q * EP;
            int64_t nb = (l < x.K) ? (x.N - col) * (int64_t) sizeof(T) : 0;
            nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
            const T* src = nb > 0 ? x.X + l * x.ldx + col : x.X;
            cp_async16(st + (size_t) p * 16, src, (uint32_t) nb);
        }
        const uint8_t* esrc = reinterpret_cast<const uint8_t*>(a.ent + (size_t) ch * x.ent_stride);
        for (int p = threadIdx.x; p < ent_bytes / 16; p += SASO_THREADS) cp_async16(st + A_BYTES + p * 16, esrc + p * 16, 16);
        const uint8_t* osrc = reinterpret_cast<const uint8_t*>(a.off + (size_t) ch * x.off_stride);
        for (int p = threadIdx.x; p < off_bytes / 16; p += SASO_THREADS) cp_async16(st + A_BYTES + ent_bytes + p * 16, osrc + p * 16, 16);
    };

    // warp owns rows [bucket*32, bucket*32+32), lane owns column jb*32 + lane
    const int bucket = rb * SASO_WARPS + warp;
    const bool active = bucket < a.nbuckets;
    T acc_lo[16], acc_hi[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { acc_lo[i] = (T) 0; acc_hi[i] = (T) 0; }

    for (int s = 0; s < SASO_NST - 1; ++s) {
        if (cb + s < ce) load_stage(cb + s, s);
        cp_async_commit();
    }
    for (int ch = cb, it = 0; ch < ce; ++ch, ++it) {
        cp_async_wait<SASO_NST - 2>();
        __syncthreads();                                   // chunk `it` has landed; everyone is done with chunk it-1
        if (ch + SASO_NST - 1 < ce) load_stage(ch + SASO_NST - 1, (it + SASO_NST - 1) % SASO_NST);
        cp_async_commit();
        if (!active || a.dbg_skip == 1) continue;
        const uint8_t* st8 = smem + (size_t)(it % SASO_NST) * stage_bytes;
        const T* st = reinterpret_cast<const T*>(st8) + lane;
        const uint16_t* ents = reinterpret_cast<const uint16_t*>(st8 + A_BYTES);
        const uint32_t* offs = reinterpret_cast<const uint32_t*>(st8 + A_BYTES + ent_bytes);
        const uint32_t o0 = offs[bucket], o1 = offs[bucket + 1];
        // four gathers in flight per step; slots beyond o1 are neutralised (valid bit clear => +0.0 into row 0)
        for (uint32_t i = o0; i < o1; i += 4) {
            uint32_t e[4]; T v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { e[q] = (i + q < o1) ? ents[i + q] : 0u; v[q] = st[(e[q] & 127u) * SASO_W]; }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const T w = (e[q] & 0x2000u) ? flip_sign<T>(v[q], (e[q] >> 12) & 1u) : (T) 0;
                if (a.dbg_skip == 2) acc_lo[0] += w; else
                if (e[q] & 0x800u) acc_select16(acc_hi, (e[q] >> 7) & 15u, w); else acc_select16(acc_lo, (e[q] >> 7) & 15u, w);
            }
        }
    }
    cp_async_wait<0>();
    if (active) {
        const int64_t col = j0 + lane;
#pragma unroll
        for (int i = 0; i < SASO_RW; ++i) {
            const T acc_i = i < 16 ? acc_lo[i & 15] : acc_hi[i & 15];
            const int64_t row = (int64_t) bucket * SASO_RW + i;
            if (row < a.M && col < a.N) {
                if (a.msplit == 1) {
                    T* c = a.C + row * a.sc_i + col * a.sc_j;
                    *c = (a.beta == (T) 0) ? a.alpha * acc_i : a.alpha * acc_i + a.beta * (*c);
                } else {
                    a.partial[((size_t) ms * a.M + row) * ((size_t) a.njb * SASO_W) + col] = acc_i;
                }
            }
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) saso_reduce_kernel(const T* __restrict__ partial, int msplit, int64_t M, int64_t N, int64_t Np,
                                                         T alpha, T beta, T* __restrict__ C, int64_t sc_i, int64_t sc_j) {
    const int64_t total = M * N;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        const int64_t row = e / N, col = e - row * N;
        T s = (T) 0;
        for (int p = 0; p < msplit; ++p) s += partial[((size_t) p * M + row) * Np + col];
        T* c = C + row * sc_i + col * sc_j;
        *c = (beta == (T) 0) ? alpha * s : alpha * s + beta * (*c);
    }
}

template <typename T>
int launch_saso_fast(const Ctr128& ctr, const Key64& key, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, int64_t ro, int64_t co,
                     int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                     T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    (void) S_cols;
    // envelope of this kernel; everything else goes to the general path
    if (sx_j != 1 || vec_nnz > 32 || M > 4096 || M < 1 || N < 1 || K < SASO_MC) return RBX_ERR_UNSUPPORTED;
    if (reinterpret_cast<uintptr_t>(X) % 16 != 0 || (sx_l * (int64_t) sizeof(T)) % 16 != 0) return RBX_ERR_UNSUPPORTED;
    if (K > (int64_t) 0x7fffffff - SASO_MC || N > 0x7fffffff) return RBX_ERR_UNSUPPORTED;
    if (alpha == (T) 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);

    const int nbuckets = (int)((M + SASO_RW - 1) / SASO_RW);
    const int nrb = (nbuckets + SASO_WARPS - 1) / SASO_WARPS;           // row blocks of 512 rows
    const int njb = (int)((N + SASO_W - 1) / SASO_W);
    const int nchunks = (int)((K + SASO_MC - 1) / SASO_MC);
    // split the contraction axis so that the grid is several waves of CTAs (one CTA per SM)
    int msplit = (6 * kNumSMs + njb * nrb - 1) / (njb * nrb);
    if (msplit > nchunks / 16) msplit = nchunks / 16;
    if (msplit < 1) msplit = 1;
    if (const char* ev = getenv("RBX_SASO_MSPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= nchunks) msplit = v; }

    const int ent_stride = SASO_MC * (int) vec_nnz;                     // multiple of 128 codes => 16-byte multiple
    const int off_stride = (nbuckets + 1 + 3) & ~3;
    uint16_t* ent = nullptr; uint32_t* off = nullptr; T* partial = nullptr;
    const size_t ent_bytes = (size_t) nchunks * ent_stride * sizeof(uint16_t);
    const size_t off_bytes = (size_t) nchunks * off_stride * sizeof(uint32_t);
    const size_t Np = (size_t) njb * SASO_W;
    const size_t part_bytes = msplit > 1 ? (size_t) msplit * M * Np * sizeof(T) : 0;
    if (cudaMallocAsync(&ent, ent_bytes, stream) != cudaSuccess || cudaMallocAsync(&off, off_bytes, stream) != cudaSuccess ||
        (part_bytes && cudaMallocAsync(&partial, part_bytes, stream) != cudaSuccess)) {
        cudaGetLastError();
        if (ent) cudaFreeAsync(ent, stream);
        if (off) cudaFreeAsync(off, stream);
        return RBX_ERR_ALLOC;
    }
    int rc = RBX_OK;
    {
        const size_t sh = (size_t) SASO_MC * vec_nnz * 2 * sizeof(uint16_t);
        if (vec_nnz <= 8) saso_prep_kernel<8><<<nchunks, SASO_MC, sh, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, nbuckets, off_stride, ent, off);
        else              saso_prep_kernel<32><<<nchunks, SASO_MC, sh, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, nbuckets, off_stride, ent, off);
        count_launch();
        rc = check_launch("saso_prep_kernel");
    }
    if (rc == RBX_OK) {
        SasoArgs<T> a;
        a.ent = ent; a.off = off; a.vec_nnz = (int) vec_nnz; a.nbuckets = nbuckets; a.njb = njb; a.msplit = msplit; a.nchunks = nchunks;
        a.nrb = nrb; a.dbg_skip = getenv("RBX_SASO_SKIP") ? atoi(getenv("RBX_SASO_SKIP")) : 0;
        a.M = M; a.N = N; a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j; a.partial = partial;
        SasoStream<T> x;
        x.X = X; x.ldx = sx_l; x.K = K; x.N = N; x.ent_stride = ent_stride; x.off_stride = off_stride;
        const size_t smem = (size_t) SASO_NST * (SASO_MC * SASO_W * sizeof(T) + ent_stride * 2 + off_stride * 4);
        if (smem > 227 * 1024) rc = RBX_ERR_UNSUPPORTED;
        if (rc == RBX_OK) {
            cudaFuncSetAttribute(saso_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
            saso_rows_kernel<T><<<(unsigned)(njb * msplit * nrb), SASO_THREADS, smem, stream>>>(x, a);
            count_launch();
            rc = check_launch("saso_rows_kernel");
        }
        if (rc == RBX_OK && msplit > 1) {
            int64_t blocks = (M * N + 255) / 256;
            if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
            saso_reduce_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(partial, msplit, M, N, (int64_t) Np, alpha, beta, C, sc_i, sc_j);
            count_launch();
            rc = check_launch("saso_reduce_kernel");
        }
    }
    cudaFreeAsync(ent, stream); cudaFreeAsync(off, stream);
    if (partial) cudaFreeAsync(partial, stream);
    return rc;
}

template int launch_saso_fast<float>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, float,
                                     const float*, int64_t, int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_saso_fast<double>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, double,
                                      const double*, int64_t, int64_t, double, double*, int64_t, int64_t, cudaStream_t);

}  // namespace rbx
