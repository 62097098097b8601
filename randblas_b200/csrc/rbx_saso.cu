// K4: streaming SASO sketch  B (d x n) = alpha * S (d x m) * A (m x n) + beta * B,  S = wide Short-axis sparse operator
// with vec_nnz signed ones per column (the lskges hot case, BASELINE config 2), A in either layout.
//
// Replaces sparse::lskges -> left_spmm -> coo::apply_coo_left_jki_p11 (skge.hh:485-510, spmm_dispatch.hh:48-160,
// coo_spmm_impl.hh:79-162): the reference sorts the triplets on the host and makes n passes over them.
//
// Design (what was measured on B200 is in DESIGN.md, "SASO kernel"):
//   * a CTA owns a tile of B, 1024 rows x one 128-byte row segment (16 doubles / 32 floats), resident in SHARED memory
//     for the whole pass; warp w owns 64 of those rows, so updates need no atomics and no inter-warp ordering;
//   * A streams through a shared-memory ring in 128-row chunks: cp.async for row-contiguous data, or vector loads + an
//     XOR-swizzled transposing store for column-contiguous data (the ring layout is the same for both);
//   * an update is: broadcast-load one 16-bit code, gather 128 B of the A chunk, read-modify-write 128 B of the B tile,
//     all conflict-free. A double-precision warp instruction carries two entries (one per half-warp);
//   * the operator is regenerated on the device (bit-exact Fisher-Yates, rbx_fy.cuh) by a prep kernel that buckets the
//     entries of each 128-column chunk of S by owning warp, in a fixed order => run-to-run identical results;
//   * the row blocks (d / 1024) of one column segment are adjacent CTAs and re-read the A chunk through L2, so HBM
//     sees each byte of A once.
// Rejected after measurement: TMA boxes for the A chunk (256-byte box rows are TMA row-rate bound, ~28 cycles/row),
// cluster multicast (73 cycles/row), register-resident tiles selected with brx.idx (~45 issue slots per select).
// Bound: shared-memory bandwidth (24 B per update) and the L2->SM re-read, not HBM; see DESIGN.md for the arithmetic.
#include <cstdlib>
#include "rbx_kernels.cuh"
#include "rbx_fy.cuh"
#include "rbx_tma.cuh"

namespace rbx {

constexpr int SASO_MC = 128;        // rows of A per chunk (= columns of S per chunk)
constexpr int SASO_ROWB = 128;      // bytes of one row segment of A / B handled by a CTA
constexpr int SASO_WARPS = 16;
constexpr int SASO_RW = 64;         // rows of B per warp
constexpr int SASO_R = SASO_WARPS * SASO_RW;   // rows of B per CTA
constexpr int SASO_THREADS = SASO_WARPS * 32;
constexpr int SASO_MAXST = 4;       // ring stages (3 when vec_nnz is large)

// ---- prep: bucket the entries of each chunk by owning warp ------------------------------------------------
// ent[chunk * MC * k + off[chunk][b] ...] = 16-bit codes: c_local (7) | r_local (6) << 7 | negative << 13 | valid << 14,
// in (column, draw) order inside each bucket.
template <int KMAX>
__global__ void __launch_bounds__(SASO_MC) saso_prep_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t ro, int64_t co,
                                                          int64_t M, int64_t K, int nbuckets, int off_stride, uint16_t* __restrict__ ent,
                                                          uint32_t* __restrict__ off) {
    extern __shared__ uint16_t sh[];
    uint16_t* code = sh;                               // MC * vec_nnz
    uint16_t* bkt = sh + SASO_MC * vec_nnz;            // MC * vec_nnz
    __shared__ uint32_t cnt[SASO_MC + 1];
    const int t = threadIdx.x;
    const int64_t chunk = blockIdx.x;
    const int64_t l = chunk * SASO_MC + t;
    for (int j = 0; j < vec_nnz; ++j) bkt[t * vec_nnz + j] = 0xFFFF;
    if (l < K) {
        fy_vector<KMAX>(ctr, key, co + l, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
            const int64_t r = idx - ro;
            if (r >= 0 && r < M) {
                code[t * vec_nnz + j] = (uint16_t)(t | ((int)(r % SASO_RW) << 7) | (negative ? 1 << 13 : 0) | (1 << 14));
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
    for (int b = t; b < off_stride; b += SASO_MC) o[b] = cnt[b <= nbuckets ? b : nbuckets];
    if (t < nbuckets) {
        uint16_t* dst = ent + chunk * total + cnt[t];
        for (int e = 0; e < total; ++e)
            if (bkt[e] == t) *dst++ = code[e];
    }
}

template <typename T>
struct SasoArgs {
    const T* X; int64_t sx_l, sx_j; int64_t K, N;      // data view: X[l * sx_l + j * sx_j], one of the strides is 1
    const uint16_t* ent; const uint32_t* off;
    int ent_stride, off_stride;                        // codes / uint32 per chunk
    int nbuckets, njb, nrb, msplit, nchunks, nst, dbg_skip;
    int64_t M;
    T alpha, beta;
    T* C; int64_t sc_i, sc_j;
    T* partial;                                        // [msplit][M][njb * W] when msplit > 1
};

template <typename T> __device__ __forceinline__ T flip_sign(T v, uint32_t neg);
template <> __device__ __forceinline__ double flip_sign<double>(double v, uint32_t neg) {
    return __hiloint2double(__double2hiint(v) ^ (int)(neg << 31), __double2loint(v));
}
template <> __device__ __forceinline__ float flip_sign<float>(float v, uint32_t neg) {
    return __int_as_float(__float_as_int(v) ^ (int)(neg << 31));
}

// 16-byte asynchronous global->shared copy; bytes beyond src_bytes are zero-filled (ragged edges)
__device__ __forceinline__ void cp_async16(void* dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// Ring layout of one A chunk: row c is 128 bytes; its 16-byte pieces are XOR-swizzled with (c >> 1) & 7 so that both the
// gather (lanes along the row) and the transposing store of column-contiguous data (lanes along c) spread over the banks.
__device__ __forceinline__ uint32_t ring_offset(uint32_t c, uint32_t byte_in_row) {
    return c * SASO_ROWB + ((((byte_in_row >> 4) ^ (c >> 1)) & 7u) << 4) + (byte_in_row & 15u);
}

// COLS_CONTIG: columns of the data view are contiguous (sx_l == 1); otherwise rows are (sx_j == 1).
template <typename T, bool COLS_CONTIG>
__global__ void __launch_bounds__(SASO_THREADS, 1) saso_tile_kernel(const SasoArgs<T> a) {
    constexpr int W = SASO_ROWB / (int) sizeof(T);     // columns per CTA: 16 doubles / 32 floats
    constexpr int EPW = 32 / W;                        // entries per warp instruction: 2 / 1
    constexpr int EP = 16 / (int) sizeof(T);           // elements per 16-byte piece
    constexpr int A_BYTES = SASO_MC * SASO_ROWB;
    extern __shared__ __align__(128) uint8_t smem[];
    T* Bs = reinterpret_cast<T*>(smem);                                   // [SASO_R][W]
    uint8_t* stages = smem + (size_t) SASO_R * SASO_ROWB;
    const int ent_bytes = a.ent_stride * 2, off_bytes = a.off_stride * 4;
    const int stage_bytes = A_BYTES + ent_bytes + off_bytes;
    const int nst = a.nst;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sub = lane / W, jl = lane % W;
    // the row blocks of one (column segment, split) are adjacent block ids: they run together and share A through L2
    const int rb = blockIdx.x % a.nrb;
    const int cid = blockIdx.x / a.nrb;
    const int jb = cid % a.njb;
    const int ms = cid / a.njb;
    const int per = (a.nchunks + a.msplit - 1) / a.msplit;
    const int cb = ms * per;
    const int ce = (cb + per < a.nchunks) ? cb + per : a.nchunks;
    const int64_t j0 = (int64_t) jb * W;

    for (int e = tid; e < SASO_R * SASO_ROWB / 16; e += SASO_THREADS) reinterpret_cast<uint4*>(smem)[e] = make_uint4(0, 0, 0, 0);

    // operator codes + bucket offsets of chunk ch -> stage s (both layouts), and the A chunk when rows are contiguous
    auto load_async = [&](int ch, int s) {
        uint8_t* st = stages + (size_t) s * stage_bytes;
        if (!COLS_CONTIG) {
            for (int p = tid; p < SASO_MC * 8; p += SASO_THREADS) {
                const int row = p >> 3, q = p & 7;
                const int64_t l = (int64_t) ch * SASO_MC + row, col = j0 + (int64_t) q * EP;
                int64_t nb = (l < a.K) ? (a.N - col) * (int64_t) sizeof(T) : 0;
                nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
                const T* src = nb > 0 ? a.X + l * a.sx_l + col : a.X;
                cp_async16(st + ring_offset(row, q * 16), src, (uint32_t) nb);
            }
        }
        const uint8_t* esrc = reinterpret_cast<const uint8_t*>(a.ent + (size_t) ch * a.ent_stride);
        for (int p = tid; p < ent_bytes / 16; p += SASO_THREADS) cp_async16(st + A_BYTES + p * 16, esrc + p * 16, 16);
        const uint8_t* osrc = reinterpret_cast<const uint8_t*>(a.off + (size_t) ch * a.off_stride);
        for (int p = tid; p < off_bytes / 16; p += SASO_THREADS) cp_async16(st + A_BYTES + ent_bytes + p * 16, osrc + p * 16, 16);
    };
    // column-contiguous data: 16-byte vector loads down the columns into registers ...
    constexpr int PPC = SASO_MC / EP;                  // pieces per column
    constexpr int NPIECE = W * PPC / SASO_THREADS;     // = 2
    T regs[NPIECE][EP];
    auto load_cols = [&](int ch) {
#pragma unroll
        for (int i = 0; i < NPIECE; ++i) {
            const int p = tid + i * SASO_THREADS, jj = p / PPC, c0 = (p % PPC) * EP;
            const int64_t l = (int64_t) ch * SASO_MC + c0, col = j0 + jj;
            const T* src = a.X + l + col * a.sx_j;
            if (col < a.N && l + EP <= a.K) {
                const uint4 v = __ldcs(reinterpret_cast<const uint4*>(src));
                *reinterpret_cast<uint4*>(regs[i]) = v;
            } else {
#pragma unroll
                for (int q = 0; q < EP; ++q) regs[i][q] = (col < a.N && l + q < a.K) ? src[q] : (T) 0;
            }
        }
    };
    // ... and an element-wise transposing store into the swizzled ring rows
    auto store_cols = [&](int s) {
        uint8_t* st = stages + (size_t) s * stage_bytes;
#pragma unroll
        for (int i = 0; i < NPIECE; ++i) {
            const int p = tid + i * SASO_THREADS, jj = p / PPC, c0 = (p % PPC) * EP;
#pragma unroll
            for (int q = 0; q < EP; ++q) *reinterpret_cast<T*>(st + ring_offset(c0 + q, jj * (int) sizeof(T))) = regs[i][q];
        }
    };

    const int dist = COLS_CONTIG ? 2 : nst - 1;        // prefetch distance in chunks
    for (int s = 0; s < dist; ++s) {
        if (cb + s < ce) {
            load_async(cb + s, s);
            if (COLS_CONTIG) { load_cols(cb + s); store_cols(s); }
        }
        cp_async_commit();
    }

    const int bucket = rb * SASO_WARPS + warp;          // warp owns rows [bucket*64, bucket*64+64) of the window
    const bool active = bucket < a.nbuckets;
    T* bwarp = Bs + (size_t) warp * SASO_RW * W + jl;

    for (int ch = cb, it = 0; ch < ce; ++ch, ++it) {
        if (dist >= 3) cp_async_wait<2>(); else cp_async_wait<1>();
        __syncthreads();                                // chunk `it` has landed; everyone is done with chunk it-1
        const int nxt = ch + dist;
        if (nxt < ce) { load_async(nxt, (it + dist) % nst); if (COLS_CONTIG) load_cols(nxt); }
        cp_async_commit();
        if (active && a.dbg_skip != 1) {
            const uint8_t* st8 = stages + (size_t)(it % nst) * stage_bytes;
            const uint16_t* ents = reinterpret_cast<const uint16_t*>(st8 + A_BYTES);
            const uint32_t* offs = reinterpret_cast<const uint32_t*>(st8 + A_BYTES + ent_bytes);
            const uint32_t o0 = offs[bucket], o1 = offs[bucket + 1];
            for (uint32_t base = o0; base < o1; base += 2 * EPW) {
                uint32_t r[2]; T w[2]; bool ok[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const uint32_t idx = base + u * EPW + sub;
                    ok[u] = idx < o1;
                    const uint32_t e = ok[u] ? ents[idx] : 0u;
                    const T v = *reinterpret_cast<const T*>(st8 + ring_offset(e & 127u, jl * (int) sizeof(T)));
                    w[u] = ok[u] ? flip_sign<T>(v, (e >> 13) & 1u) : (T) 0;
                    r[u] = (e >> 7) & 63u;
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    T* b = bwarp + r[u] * W;
                    if (EPW == 2) {
                        // the two half-warps carry different entries; if both hit the same row they must not race
                        const uint32_t other = __shfl_xor_sync(0xffffffffu, ok[u] ? r[u] : 0xffu, 16);
                        const bool clash = ok[u] && other == r[u];
                        if (!clash) { if (ok[u]) *b += w[u]; }
                        else { if (sub == 0) *b += w[u]; __syncwarp(); if (sub == 1) *b += w[u]; __syncwarp(); }
                    } else {
                        if (ok[u]) *b += w[u];
                    }
                }
            }
        }
        if (COLS_CONTIG && nxt < ce) store_cols((it + dist) % nst);
    }
    cp_async_wait<0>();
    __syncthreads();
    if (active) {
        const int64_t col = j0 + jl;
        for (int i = sub; i < SASO_RW; i += EPW) {
            const int64_t row = (int64_t) bucket * SASO_RW + i;
            const T acc = bwarp[i * W];
            if (row < a.M && col < a.N) {
                if (a.msplit == 1) {
                    T* c = a.C + row * a.sc_i + col * a.sc_j;
                    *c = (a.beta == (T) 0) ? a.alpha * acc : a.alpha * acc + a.beta * (*c);
                } else {
                    a.partial[((size_t) ms * a.M + row) * ((size_t) a.njb * W) + col] = acc;
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
    constexpr int W = SASO_ROWB / (int) sizeof(T);
    // envelope of this kernel; everything else goes to the general path
    const bool cols_contig = (sx_l == 1 && sx_j != 1);
    if (!cols_contig && sx_j != 1) return RBX_ERR_UNSUPPORTED;
    if (vec_nnz > 32 || M > SASO_RW * SASO_MC || M < 1 || N < 1 || K < SASO_MC) return RBX_ERR_UNSUPPORTED;
    const int64_t ld = cols_contig ? sx_j : sx_l;
    if (reinterpret_cast<uintptr_t>(X) % 16 != 0 || (ld * (int64_t) sizeof(T)) % 16 != 0) return RBX_ERR_UNSUPPORTED;
    if (K > (int64_t) 0x7fffffff - SASO_MC || N > 0x7fffffff) return RBX_ERR_UNSUPPORTED;
    if (alpha == (T) 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);

    const int nbuckets = (int)((M + SASO_RW - 1) / SASO_RW);            // <= 128 == threads of the prep kernel
    const int nrb = (nbuckets + SASO_WARPS - 1) / SASO_WARPS;           // row blocks of 1024 rows
    const int njb = (int)((N + W - 1) / W);
    const int nchunks = (int)((K + SASO_MC - 1) / SASO_MC);
    // split the contraction axis so that the grid is several waves of CTAs (one CTA per SM)
    int msplit = (6 * kNumSMs + njb * nrb - 1) / (njb * nrb);
    if (msplit > nchunks / 16) msplit = nchunks / 16;
    if (msplit < 1) msplit = 1;
    if (const char* ev = getenv("RBX_SASO_MSPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= nchunks) msplit = v; }

    const int ent_stride = SASO_MC * (int) vec_nnz;                     // a multiple of 128 codes => 16-byte multiple
    const int off_stride = (nbuckets + 1 + 3) & ~3;
    const size_t stage = (size_t) SASO_MC * SASO_ROWB + ent_stride * 2 + off_stride * 4;
    int nst = SASO_MAXST;
    while (nst > 3 && (size_t) SASO_R * SASO_ROWB + nst * stage > 227 * 1024) --nst;
    const size_t smem = (size_t) SASO_R * SASO_ROWB + nst * stage;
    if (smem > 227 * 1024) return RBX_ERR_UNSUPPORTED;

    uint16_t* ent = nullptr; uint32_t* off = nullptr; T* partial = nullptr;
    const size_t ent_bytes = (size_t) nchunks * ent_stride * sizeof(uint16_t);
    const size_t off_bytes = (size_t) nchunks * off_stride * sizeof(uint32_t);
    const size_t Np = (size_t) njb * W;
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
        a.X = X; a.sx_l = sx_l; a.sx_j = sx_j; a.K = K; a.N = N;
        a.ent = ent; a.off = off; a.ent_stride = ent_stride; a.off_stride = off_stride;
        a.nbuckets = nbuckets; a.njb = njb; a.nrb = nrb; a.msplit = msplit; a.nchunks = nchunks; a.nst = nst;
        a.dbg_skip = getenv("RBX_SASO_SKIP") ? atoi(getenv("RBX_SASO_SKIP")) : 0;
        a.M = M; a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j; a.partial = partial;
        const unsigned grid = (unsigned)(njb * msplit * nrb);
        if (cols_contig) {
            cudaFuncSetAttribute(saso_tile_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
            saso_tile_kernel<T, true><<<grid, SASO_THREADS, smem, stream>>>(a);
        } else {
            cudaFuncSetAttribute(saso_tile_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
            saso_tile_kernel<T, false><<<grid, SASO_THREADS, smem, stream>>>(a);
        }
        count_launch();
        rc = check_launch("saso_tile_kernel");
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
