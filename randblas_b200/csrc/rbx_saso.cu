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
                                                          int64_t M, int64_t K, int nbuckets, uint16_t* __restrict__ ent,
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
    uint32_t* o = off + chunk * (nbuckets + 1);
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
    int vec_nnz, nbuckets, njb, msplit, nchunks;
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

template <typename T>
__global__ void __launch_bounds__(SASO_THREADS, 1) saso_rows_kernel(const __grid_constant__ CUtensorMap tmap, const SasoArgs<T> a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    T* stages = reinterpret_cast<T*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t) SASO_NST * SASO_MC * SASO_W * sizeof(T));
    uint64_t* empty = full + SASO_NST;

    const uint32_t cs = cluster_nctarank(), rank = cluster_ctarank();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cid = blockIdx.x / cs;
    const int jb = cid % a.njb;
    const int ms = (cid / a.njb) % a.msplit;
    const int rg = cid / (a.njb * a.msplit);
    const int per = (a.nchunks + a.msplit - 1) / a.msplit;
    const int cb = ms * per;
    const int ce = (cb + per < a.nchunks) ? cb + per : a.nchunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SASO_NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], cs * SASO_WARPS); }
        fence_mbar_init();
    }
    cluster_sync_all();

    // ===== producer duty (warp 0, lane 0): this CTA loads its 1/cs slice of every chunk into all CTAs of the cluster =====
    const int rows_per = SASO_MC / (int) cs;
    const uint16_t mask = (uint16_t)((1u << cs) - 1);
    auto issue = [&](int it) {
        const int s = it % SASO_NST;
        mbar_expect_tx(&full[s], SASO_MC * SASO_W * sizeof(T));
        T* dst = stages + (size_t) s * SASO_MC * SASO_W + (size_t) rank * rows_per * SASO_W;
        const int y = (cb + it) * SASO_MC + (int) rank * rows_per;
        if (cs > 1) tma_load_2d_mcast(dst, &tmap, jb * SASO_W, y, &full[s], mask);
        else        tma_load_2d(dst, &tmap, jb * SASO_W, y, &full[s]);
    };
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmap);
        for (int it = 0; it < SASO_NST && cb + it < ce; ++it) issue(it);
    }
    __syncwarp();
    {
        // ===== consumers: warp owns rows [bucket*32, bucket*32+32), lane owns column jb*32 + lane =====
        const int bucket = (rg * (int) cs + (int) rank) * SASO_WARPS + warp;
        const bool active = bucket < a.nbuckets;
        T acc_lo[16], acc_hi[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) { acc_lo[i] = (T) 0; acc_hi[i] = (T) 0; }
        const int per_chunk = SASO_MC * a.vec_nnz;
        uint32_t o0 = 0, o1 = 0, e_lane = 0;
        if (active && cb < ce) {
            const uint32_t* o = a.off + (size_t) cb * (a.nbuckets + 1) + bucket;
            o0 = o[0]; o1 = o[1];
            if (o0 + lane < o1) e_lane = a.ent[(size_t) cb * per_chunk + o0 + lane];
        }
        for (int ch = cb, it = 0; ch < ce; ++ch, ++it) {
            const int s = it % SASO_NST;
            const uint32_t ph = (it / SASO_NST) & 1;
            // prefetch the next chunk's first 32 codes while this one is processed
            uint32_t n0 = 0, n1 = 0, ne = 0;
            if (active && ch + 1 < ce) {
                const uint32_t* o = a.off + (size_t)(ch + 1) * (a.nbuckets + 1) + bucket;
                n0 = o[0]; n1 = o[1];
                if (n0 + lane < n1) ne = a.ent[(size_t)(ch + 1) * per_chunk + n0 + lane];
            }
            mbar_wait(&full[s], ph);
            const T* st = stages + (size_t) s * SASO_MC * SASO_W + lane;
            for (uint32_t base = o0; base < o1; base += 32) {
                if (base != o0) e_lane = (base + lane < o1) ? a.ent[(size_t) ch * per_chunk + base + lane] : 0;
                const int cnt = (o1 - base < 32u) ? (int)(o1 - base) : 32;
                // four gathers in flight per step; codes beyond cnt are 0 (valid bit clear) and add +0.0 to row 0
                for (int i = 0; i < cnt; i += 4) {
                    uint32_t e[4]; T v[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) e[q] = __shfl_sync(0xffffffffu, e_lane, (i + q) & 31);
#pragma unroll
                    for (int q = 0; q < 4; ++q) { if (i + q >= cnt) e[q] = 0; v[q] = st[(e[q] & 127u) * SASO_W]; }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const T w = (e[q] & 0x2000u) ? flip_sign<T>(v[q], (e[q] >> 12) & 1u) : (T) 0;
                        if (e[q] & 0x800u) acc_select16(acc_hi, (e[q] >> 7) & 15u, w); else acc_select16(acc_lo, (e[q] >> 7) & 15u, w);
                    }
                }
            }
            __syncwarp();
            if (lane < (int) cs) mbar_arrive_cluster(&empty[s], lane);   // release the stage in every CTA that multicasts into it
            if (threadIdx.x == 0 && ch + SASO_NST < ce) {               // refill it once every consumer of the cluster is done
                mbar_wait_cluster(&empty[s], ph);
                issue(it + SASO_NST);
            }
            __syncwarp();
            o0 = n0; o1 = n1; e_lane = ne;
        }
        if (active) {
            const int64_t col = (int64_t) jb * SASO_W + lane;
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
    cluster_sync_all();   // no CTA may exit while a sibling can still multicast into it or arrive on its barriers
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

template <typename T> struct TmapType;
template <> struct TmapType<float>  { static constexpr CUtensorMapDataType v = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; };
template <> struct TmapType<double> { static constexpr CUtensorMapDataType v = CU_TENSOR_MAP_DATA_TYPE_FLOAT64; };

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
    int cs = 1; while (cs < nrb && cs < 8) cs *= 2;                     // cluster = row blocks sharing the A stream
    const int nrg = (nrb + cs - 1) / cs;
    const int njb = (int)((N + SASO_W - 1) / SASO_W);
    const int nchunks = (int)((K + SASO_MC - 1) / SASO_MC);
    // split the contraction axis so that the grid is several waves of co-resident clusters
    const int conc = kNumSMs / cs;
    int msplit = (4 * conc + njb * nrg - 1) / (njb * nrg);
    if (msplit > nchunks / 16) msplit = nchunks / 16;
    if (msplit < 1) msplit = 1;

    uint16_t* ent = nullptr; uint32_t* off = nullptr; T* partial = nullptr;
    const size_t ent_bytes = (size_t) nchunks * SASO_MC * vec_nnz * sizeof(uint16_t);
    const size_t off_bytes = (size_t) nchunks * (nbuckets + 1) * sizeof(uint32_t);
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
        if (vec_nnz <= 8) saso_prep_kernel<8><<<nchunks, SASO_MC, sh, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, nbuckets, ent, off);
        else              saso_prep_kernel<32><<<nchunks, SASO_MC, sh, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, nbuckets, ent, off);
        count_launch();
        rc = check_launch("saso_prep_kernel");
    }
    CUtensorMap tmap;
    if (rc == RBX_OK && !make_tmap_2d(&tmap, TmapType<T>::v, X, (uint64_t) N, (uint64_t) K, (uint64_t) sx_l * sizeof(T), SASO_W, SASO_MC / cs))
        rc = RBX_ERR_UNSUPPORTED;
    if (rc == RBX_OK) {
        SasoArgs<T> a;
        a.ent = ent; a.off = off; a.vec_nnz = (int) vec_nnz; a.nbuckets = nbuckets; a.njb = njb; a.msplit = msplit; a.nchunks = nchunks;
        a.M = M; a.N = N; a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j; a.partial = partial;
        const size_t smem = (size_t) SASO_NST * SASO_MC * SASO_W * sizeof(T) + 2 * SASO_NST * sizeof(uint64_t);
        static bool attr_set[2] = { false, false };
        const int ti = sizeof(T) == 8 ? 1 : 0;
        if (!attr_set[ti]) {
            cudaFuncSetAttribute(saso_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
            attr_set[ti] = true;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(njb * msplit * nrg * cs));
        cfg.blockDim = dim3(SASO_THREADS);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, saso_rows_kernel<T>, tmap, a) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
        count_launch();
        if (rc == RBX_OK) rc = check_launch("saso_rows_kernel");
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
