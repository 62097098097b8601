// K2 (tensor-core path, float): fused generate-and-contract GEMM on the 5th-generation tensor cores.
//
//   C (M x N) = alpha * Sop (M x K) * X (K x N) + beta * C,   Sop = a window of the implicit dense operator
//
// Replaces dense::lskge3 / rskge3 (skge.hh:173-215, 320-364) for large aligned float problems (BASELINE config 1):
// the reference writes the d x m operator to memory and calls sgemm; here the operator tile goes from the Philox
// counter straight into the shared-memory operand of tcgen05.mma and never exists in HBM.
//
//  * 128 x BN output tile per CTA (BN = 256: the FP32 accumulator is 128 lanes x 256 columns of TMEM), BK = 32 floats
//    = one 128-byte swizzle atom per row; rings: 3 X tiles (TMA latency), 2 operator tiles and 2 X_lo tiles (224 KB).
//  * FP32 accuracy from TF32 tensor cores: 3xTF32. Both operands are split x = hi + lo (hi = top 19 bits, lo = x - hi,
//    exact in FP32) and D += hi*hi + lo*hi + hi*lo with FP32 accumulation in TMEM; the dropped lo*lo term is 2^-22
//    relative, inside the reference's own GEMM tolerance (test/comparison.hh).
//  * warp roles: warp 0 = TMA producer for X (cp.async.bulk.tensor, 128B swizzle), warp 1 = TMEM allocator + single
//    MMA-issuing thread, warps 2..17 = generators: bit-exact Philox/Box-Muller (rbx_rng.cuh) -> hi/lo split -> swizzled
//    K-major operand layout, then the hi/lo split of the X tile in place; they are also the epilogue warps
//    (tcgen05.ld -> alpha/beta -> global, or split-K partials summed in a fixed order by a second kernel).
//  * the CTAs of one row block (same operator rows, different columns of X) form a thread-block cluster of up to 4:
//    each generates 1/CL of the operator tile into its own operand ring (plain vector stores) and a forwarder warp sends
//    those rows to every peer with one bulk shared-to-shared copy per tile half (cp.async.bulk.shared::cluster, completes
//    bytes on the peer's mbarrier), so a Gaussian sample is generated once per cluster;
//  * mbarriers: full_x (TMA -> generators), gen_done (own rows stored -> MMA + forwarder), ready_s (bulk-copy bytes from the peers -> MMA), ready_x (generators ->
//    MMA), empty (tcgen05.commit multicast to the whole cluster -> TMA + generators), done (last commit -> epilogue).
// Envelope: implicit operator whose parent rows run along the contraction axis (l_on_cols), window column offset a
// multiple of 4, X with unit stride along the contraction axis (e.g. ColMajor A in a left sketch), 16-byte aligned.
// Everything else returns RBX_ERR_UNSUPPORTED and takes the SIMT kernel.
#include <cstdlib>
#include "rbx_kernels.cuh"
#include "rbx_copy.cuh"
#include "rbx_tma.cuh"

namespace rbx {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 32;
#ifndef RBX_TC_MIN_N
#define RBX_TC_MIN_N 2        // tall-skinny data: the padded 256-column tile still beats the SIMT kernel 17-27x (measured at n = 4 .. 48)
#endif
#ifndef RBX_TC_MIN_K
#define RBX_TC_MIN_K 32       // a short contraction axis ("lifts", B = S A with a tall S): the kernel is bound by its epilogue there and still 2-5x ahead of the SIMT kernel
#endif
#ifndef TC_X_HI_RAW
#define TC_X_HI_RAW 1
#endif
constexpr int TC_XST = 3;           // raw / hi X tiles in flight (hides the TMA latency)
constexpr int TC_LST = 2;           // operator (hi + lo) tiles and X_lo tiles (produced on chip, one stage ahead is enough)
constexpr int TC_EPI_WARPS = 16;     // generator warps that also run the epilogue (a multiple of 4)
constexpr int TC_GEN_WARPS = 16, TC_THREADS = (3 + TC_GEN_WARPS) * 32;   // TMA, MMA, 16 generators, forwarder
constexpr int TC_S_BYTES = TC_BM * TC_BK * 4;          // 16 KB: one of S_hi / S_lo
constexpr int TC_X_BYTES = TC_BN * TC_BK * 4;          // 32 KB: one of X_hi / X_lo
// shared-memory map: [X hi x3][X lo x2][S hi, S lo x2]
constexpr int TC_OFF_XLO = TC_XST * TC_X_BYTES;
constexpr int TC_OFF_S = TC_OFF_XLO + TC_LST * TC_X_BYTES;
constexpr size_t TC_SMEM = (size_t) TC_OFF_S + TC_LST * 2 * TC_S_BYTES + 1024 /* alignment slack */;

// ---- tcgen05 wrappers --------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_commit_mcast(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 :: "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t cta) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major operand tile, 128-byte rows, SWIZZLE_128B: 8-row groups are 1024 bytes apart (cute/arch/mma_sm100_desc.hpp fields)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);          // start address, bits [0,14)
    d |= (uint64_t) 1 << 16;                               // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                      // stride byte offset, bits [32,46)
    d |= (uint64_t) 1 << 46;                               // descriptor version (Blackwell)
    d |= (uint64_t) 2 << 61;                               // layout type SWIZZLE_128B
    return d;
}
// kind::tf32, FP32 accumulate, both operands K-major, M = 128, N = BN
constexpr uint32_t tc_idesc(int bn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}

struct TcArgs {
    DenseOpWindow w;
    int64_t M, N, K;
    float alpha, beta;
    float* C; int64_t sc_i, sc_j;
    float* partial;               // [ksplit][Npad][Mpad] when ksplit > 1
    int64_t Mpad, Npad;
    int nk, ksplit, tiles_n, dbg, cl;
    const float* Sx; int64_t s_pitch;   // FAMILY 'X': operator values read from a k-contiguous array (panel workspace or S.buff)
};

template <char FAMILY>
__global__ void __launch_bounds__(TC_THREADS, 1) sketch_dense_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const TcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_x[TC_XST], x_empty[TC_XST], ready_s[TC_LST], ready_x[TC_LST], gen_done[TC_LST], l_empty[TC_LST], done;
    __shared__ uint32_t tmem_base_s;
    __shared__ double tab[32];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t) 1023);
    const uint32_t smem0 = smem_u32(smem);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    const int tile = blockIdx.x % (a.tiles_n * (int)((a.M + TC_BM - 1) / TC_BM));
    const int ks = blockIdx.x / (a.tiles_n * (int)((a.M + TC_BM - 1) / TC_BM));
    const int64_t i0 = (int64_t)(tile / a.tiles_n) * TC_BM, j0 = (int64_t)(tile % a.tiles_n) * TC_BN;
    const int per = (a.nk + a.ksplit - 1) / a.ksplit;
    const int kb = ks * per, ke = (kb + per < a.nk) ? kb + per : a.nk;
    const int nit = ke - kb;

    if (tid == 0) {
        for (int s = 0; s < TC_XST; ++s) { mbar_init(&full_x[s], 1); mbar_init(&x_empty[s], 1); }
        for (int s = 0; s < TC_LST; ++s) { mbar_init(&ready_s[s], 1); mbar_init(&ready_x[s], TC_GEN_WARPS); mbar_init(&gen_done[s], TC_GEN_WARPS); mbar_init(&l_empty[s], (uint32_t) a.cl); }
        mbar_init(&done, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmap_x);
    }
    if (FAMILY == 'G') load_logf_tab(tab);
    if (warp == 1) tmem_alloc(&tmem_base_s, TC_BN);
    tc_fence_before();
    __syncthreads();
    if (a.cl > 1) cluster_sync_all();                      // peers' barriers are initialised before any remote traffic
    tc_fence_after();
    const uint32_t tmem_d = tmem_base_s;
    const uint32_t crank = a.cl > 1 ? cluster_ctarank() : 0u;

    if (warp == 0) {
        // ---- TMA producer: raw X tile (BN rows of 32 floats) into the X_hi slot of the stage ----
        if (lane == 0) {
            for (int it = 0; it < nit; ++it) {
                const int s = it % TC_XST, u = it / TC_XST;
                if (u > 0) mbar_wait(&x_empty[s], (u - 1) & 1);
                if (a.dbg & 8) { mbar_arrive(&full_x[s]); continue; }
                mbar_expect_tx(&full_x[s], TC_X_BYTES);
                tma_load_2d(smem + (size_t) s * TC_X_BYTES, &tmap_x, (kb + it) * TC_BK, (int) j0, &full_x[s]);
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer ----
        if (lane == 0) {
            constexpr uint32_t idesc = tc_idesc(TC_BN);
            for (int it = 0; it < nit; ++it) {
                const int s = it % TC_LST, u = it / TC_LST, xs = it % TC_XST;
                mbar_wait(&gen_done[s], u & 1);            // this CTA's rows of the operator tile (generic stores + proxy fence)
                mbar_wait(&ready_s[s], u & 1);             // the peers' rows (bulk copies into this CTA)
                mbar_wait(&ready_x[s], u & 1);
                fence_proxy_async();                       // st.async / generic writes -> tensor-core (async proxy) reads
                tc_fence_after();
                const uint32_t sS = smem0 + TC_OFF_S + (uint32_t) s * 2 * TC_S_BYTES;
                const uint64_t a_hi = umma_desc_sw128(sS), a_lo = umma_desc_sw128(sS + TC_S_BYTES);
                const uint64_t b_hi = umma_desc_sw128(smem0 + (uint32_t) xs * TC_X_BYTES), b_lo = umma_desc_sw128(smem0 + TC_OFF_XLO + (uint32_t) s * TC_X_BYTES);
#pragma unroll
                for (int k = 0; k < ((a.dbg & 4) ? 0 : TC_BK / 8); ++k) {
                    const uint64_t adv = (uint64_t)(k * 8 * 4 >> 4);          // 32 bytes per K step inside the swizzle atom
                    umma_tf32(tmem_d, a_lo + adv, b_hi + adv, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    umma_tf32(tmem_d, a_hi + adv, b_lo + adv, idesc, 1u);
                    umma_tf32(tmem_d, a_hi + adv, b_hi + adv, idesc, 1u);
                }
                // when these MMAs have read their operands: the X slot is free for the TMA producer, the operator / X_lo slot
                // is free in every CTA of the cluster (peers write operator rows into it)
                umma_commit(&x_empty[xs]);
                if (a.cl > 1) umma_commit_mcast(&l_empty[s], (uint16_t)((1u << a.cl) - 1u)); else umma_commit(&l_empty[s]);
            }
            umma_commit(&done);
        }
    } else if (warp == 2 + TC_GEN_WARPS) {
        // ---- forwarder: when this CTA's rows of a stage are in its own shared memory, one bulk copy per peer and half (hi / lo)
        // carries them into the peers' operand rings and completes the bytes on their ready_s barriers. (Round 1 handed every
        // 16-byte chunk over with its own st.async; those small async-proxy writes did not overlap with the tensor core's operand
        // reads: a probe with the stores on cost 0.76 ms beside the MMAs, 0.50 ms with them off.)
        if (lane == 0 && a.cl > 1) {
            const uint32_t part = (uint32_t)(TC_BM / a.cl) * 128u;      // bytes of this CTA's rows in one of the hi / lo tiles
            for (int it = 0; it < nit; ++it) {
                const int s = it % TC_LST, u = it / TC_LST;
                mbar_wait(&gen_done[s], u & 1);
                const uint32_t src = smem0 + TC_OFF_S + (uint32_t) s * 2 * TC_S_BYTES + crank * part, bar_u32 = smem_u32(&ready_s[s]);
                for (uint32_t p = 0; p < (uint32_t) a.cl; ++p) {
                    if (p == crank) continue;
                    const uint32_t dst = mapa_u32(src, p), bar = mapa_u32(bar_u32, p);
                    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst), "r"(src), "r"(part), "r"(bar) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst + TC_S_BYTES), "r"(src + TC_S_BYTES), "r"(part), "r"(bar) : "memory");
                }
            }
        }
    } else {
        // ---- generators / converters ----
        const int gt = tid - 64;                           // 0 .. 32 * TC_GEN_WARPS - 1
        const float scale = (FAMILY == 'U') ? (float) 1.7320508075688772 : 1.0f;
        // The operator values of stage it+1 are generated into registers while the tensor core works on stage it (after this
        // warp has handed stage it over), so the FP64-heavy Box-Muller never waits for a shared-memory slot.
        constexpr int NCH = (TC_BM * (TC_BK / 4) + TC_GEN_WARPS * 32 - 1) / (TC_GEN_WARPS * 32);
        const int rows_per = TC_BM / a.cl;                 // operator rows this CTA generates for the cluster
        const int my_chunks = (a.dbg & 1) ? 0 : rows_per * (TC_BK / 4);
        float4 vals[NCH];
        auto generate = [&](int it) {
            const int64_t l0 = (int64_t)(kb + it) * TC_BK;
            const int64_t cbase = (a.w.pc0 + l0) >> 2;     // first counter column of this K block (pc0 % 4 == 0)
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
                const int e = gt + q * TC_GEN_WARPS * 32;
                const int r = (int) crank * rows_per + (e >> 3), c = e & 7;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (a.dbg & 48) {
                    // pipe-conflict probe (tools/tc_time.py, RBX_TC_DBG 16 / 32): a dependent chain of 200 FP64 (16) or FP32 (32) FMAs
                    // per operator chunk instead of the generator, to see which of them can run beside the tensor core
                    if (a.dbg & 16) { double x = (double) e; for (int z = 0; z < 200; ++z) x = fma(x, 1.0000001, 1e-9); v.x = (float) x; }
                    else { float x = (float) e; for (int z = 0; z < 200; ++z) x = fmaf(x, 1.0000001f, 1e-9f); v.x = x; }
                } else if (e < my_chunks && i0 + r < a.M && l0 + 4 * c < a.K) {
                    if (FAMILY == 'X') {
                        v = __ldg(reinterpret_cast<const float4*>(a.Sx + (i0 + r) * a.s_pitch + l0 + 4 * c));
                    } else {
                        const Ctr128 ctr = ctr_add(a.w.ctr, (uint64_t)((a.w.pr0 + i0 + r) * a.w.W + cbase + c));
                        v = generate4<FAMILY>(philox4x32_10(ctr, a.w.key), tab);
                    }
                    if (FAMILY == 'U') { v.x = __fmul_rn(v.x, scale); v.y = __fmul_rn(v.y, scale); v.z = __fmul_rn(v.z, scale); v.w = __fmul_rn(v.w, scale); }
                    if (FAMILY != 'X') {           // 'X': masked where the values are consumed, so the load stays in flight
                        if (l0 + 4 * c + 1 >= a.K) v.y = 0.f;
                        if (l0 + 4 * c + 2 >= a.K) v.z = 0.f;
                        if (l0 + 4 * c + 3 >= a.K) v.w = 0.f;
                    }
                }
                vals[q] = v;
            }
        };
        if (nit > 0) generate(0);
        for (int it = 0; it < nit; ++it) {
            const int s = it % TC_LST, u = it / TC_LST, xs = it % TC_XST;
            if (u > 0) mbar_wait(&l_empty[s], (u - 1) & 1);
            if (gt == 0) mbar_expect_tx(&ready_s[s], (uint32_t)(2 * TC_S_BYTES / a.cl * (a.cl - 1)));      // the peers' rows of the hi + lo tiles
            uint8_t* st_loc = smem + TC_OFF_S + (size_t) s * 2 * TC_S_BYTES;
            const uint32_t bar_u32 = smem_u32(&ready_s[s]);
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
                const int e = gt + q * TC_GEN_WARPS * 32;
                if (e < my_chunks) {
                    const int r = (int) crank * rows_per + (e >> 3), c = e & 7;
                    float4 v = vals[q];
                    if (FAMILY == 'X') {
                        const int64_t lc = (int64_t)(kb + it) * TC_BK + 4 * c;
                        if (lc + 1 >= a.K) v.y = 0.f;
                        if (lc + 2 >= a.K) v.z = 0.f;
                        if (lc + 3 >= a.K) v.w = 0.f;
                    }
                    float4 hi, lo;
                    hi.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); lo.x = v.x - hi.x;
                    hi.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u); lo.y = v.y - hi.y;
                    hi.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); lo.z = v.z - hi.z;
                    hi.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u); lo.w = v.w - hi.w;
                    const uint32_t off = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
                    *reinterpret_cast<float4*>(st_loc + off) = hi;                        // this CTA's own ring; the forwarder copies the
                    *reinterpret_cast<float4*>(st_loc + TC_S_BYTES + off) = lo;           // rows to the peers in bulk
                }
            }
            fence_proxy_async();                           // generic stores -> visible to the tensor core and to the bulk-copy engine
            __syncwarp();
            if (lane == 0) mbar_arrive(&gen_done[s]);
            // operator values read from memory: issue the loads of the next stage now, so that their latency is covered by
            // the X split below (the generated families compute theirs after handing the stage over instead)
            if (FAMILY == 'X' && it + 1 < nit) generate(it + 1);
            mbar_wait(&full_x[xs], (it / TC_XST) & 1);
            float4* xh = reinterpret_cast<float4*>(smem + (size_t) xs * TC_X_BYTES);
            float4* xl = reinterpret_cast<float4*>(smem + TC_OFF_XLO + (size_t) s * TC_X_BYTES);
#pragma unroll 4
            for (int e = gt; e < ((a.dbg & 2) ? 0 : TC_X_BYTES / 16); e += TC_GEN_WARPS * 32) {
                const float4 v = xh[e];
                float4 hi, lo;
                hi.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); lo.x = v.x - hi.x;
                hi.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u); lo.y = v.y - hi.y;
                hi.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); lo.z = v.z - hi.z;
                hi.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u); lo.w = v.w - hi.w;
#if TC_X_HI_RAW
                xl[e] = lo;          // the raw tile serves as X_hi: kind::tf32 reads the top 19 bits of each operand word and ignores the rest
#else
                xh[e] = hi; xl[e] = lo;
#endif
            }
            if (!(a.dbg & 64)) fence_proxy_async();        // generic-proxy writes -> visible to the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) mbar_arrive(&ready_x[s]);
            if (FAMILY != 'X' && it + 1 < nit) generate(it + 1);
        }
        // ---- epilogue: TMEM -> registers -> global ----
        mbar_wait(&done, 0);
        tc_fence_after();
        const int lg = warp & 3;                           // TMEM lane quarter this warp may read
        constexpr int CG = TC_BN / (TC_EPI_WARPS / 4);     // columns per epilogue warp
        const int cgrp = (warp - 2) >> 2;                  // which group of the BN columns
        const int64_t row = i0 + lg * 32 + lane;
        const bool vec_rows = a.ksplit == 1 && a.sc_j == 1 && (a.sc_i & 3) == 0 && (reinterpret_cast<uintptr_t>(a.C) & 15) == 0;
        for (int c0 = cgrp * CG; c0 < (cgrp + 1) * CG && warp - 2 < TC_EPI_WARPS; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t) c0, v);
            if (row < a.M && vec_rows && j0 + c0 + 32 <= a.N) {
                // C with unit stride along j: the thread's 32 consecutive columns of its row go out as eight 16-byte stores
                float4* cp = reinterpret_cast<float4*>(a.C + row * a.sc_i + j0 + c0);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    float4 o;
                    o.x = a.alpha * ((nit > 0) ? __uint_as_float(v[4 * q + 0]) : 0.f); o.y = a.alpha * ((nit > 0) ? __uint_as_float(v[4 * q + 1]) : 0.f);
                    o.z = a.alpha * ((nit > 0) ? __uint_as_float(v[4 * q + 2]) : 0.f); o.w = a.alpha * ((nit > 0) ? __uint_as_float(v[4 * q + 3]) : 0.f);
                    if (a.beta != 0.f) { const float4 old = cp[q]; o.x += a.beta * old.x; o.y += a.beta * old.y; o.z += a.beta * old.z; o.w += a.beta * old.w; }
                    cp[q] = o;
                }
            } else if (row < a.M) {
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    const int64_t col = j0 + c0 + q;
                    if (col < a.N) {
                        const float acc = (nit > 0) ? __uint_as_float(v[q]) : 0.f;
                        if (a.ksplit == 1) {
                            float* cp = a.C + row * a.sc_i + col * a.sc_j;
                            *cp = (a.beta == 0.f) ? a.alpha * acc : a.alpha * acc + a.beta * (*cp);
                        } else {
                            a.partial[((size_t) ks * a.Npad + col) * a.Mpad + row] = acc;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (a.cl > 1) cluster_sync_all();                      // no CTA leaves while peers may still write to it or signal it
    if (warp == 1) tmem_dealloc(tmem_d, TC_BN);
}

__global__ void __launch_bounds__(256) tc_reduce_kernel(const float* __restrict__ partial, int ksplit, int64_t M, int64_t N, int64_t Mpad, int64_t Npad,
                                                      float alpha, float beta, float* __restrict__ C, int64_t sc_i, int64_t sc_j) {
    const int64_t total = M * N;
    for (int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t) gridDim.x * blockDim.x) {
        const int64_t col = e / M, row = e - col * M;
        float s = 0.f;
        for (int p = 0; p < ksplit; ++p) s += partial[((size_t) p * Npad + col) * Mpad + row];
        float* c = C + row * sc_i + col * sc_j;
        *c = (beta == 0.f) ? alpha * s : alpha * s + beta * (*c);
    }
}

// one launch of the tensor-core kernel (+ split-K reduce). mode: 'G' / 'U' = operator generated in the kernel from g.w;
// 'X' = operator values read from Sx[i * s_pitch + l].
static int tc_launch(const GemmOperands<float>& g, char mode, const float* Sx, int64_t s_pitch, cudaStream_t stream) {
    CUtensorMap tmap;
    if (!make_tmap_2d(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, g.X, (uint64_t) g.K, (uint64_t) g.N, (uint64_t) g.sx_j * 4, TC_BK, TC_BN,
                      CU_TENSOR_MAP_SWIZZLE_128B))
        return RBX_ERR_UNSUPPORTED;

    TcArgs a;
    a.Sx = Sx; a.s_pitch = s_pitch;
    a.w = g.w; a.M = g.M; a.N = g.N; a.K = g.K; a.alpha = g.alpha; a.beta = g.beta; a.C = g.C; a.sc_i = g.sc_i; a.sc_j = g.sc_j;
    const int tiles_m = (int)((g.M + TC_BM - 1) / TC_BM);
    a.tiles_n = (int)((g.N + TC_BN - 1) / TC_BN);
    a.nk = (int)((g.K + TC_BK - 1) / TC_BK);
    const int tiles = tiles_m * a.tiles_n;
    int ksplit = (2 * kNumSMs) / tiles;                    // about two waves of CTAs
    if (ksplit > a.nk / 8) ksplit = a.nk / 8;
    if (ksplit < 1) ksplit = 1;
    if (const char* ev = getenv("RBX_TC_KSPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= a.nk) ksplit = v; }
    a.ksplit = ksplit;
    a.dbg = getenv("RBX_TC_DBG") ? atoi(getenv("RBX_TC_DBG")) : 0;
    a.Mpad = (g.M + 31) & ~(int64_t) 31; a.Npad = g.N;
    a.partial = nullptr;
    if (ksplit > 1 && ws_alloc(&a.partial, sizeof(float) * (size_t) ksplit * a.Mpad * a.Npad, stream) != cudaSuccess) {
        cudaGetLastError();
        return RBX_ERR_ALLOC;
    }
    auto kern = mode == 'X' ? sketch_dense_tc_kernel<'X'> : mode == 'G' ? sketch_dense_tc_kernel<'G'> : sketch_dense_tc_kernel<'U'>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) TC_SMEM);
    // Pairs of CTAs that are neighbours along n form a cluster: each generates half of the operator tile and multicasts it
    // (st.async + tcgen05.commit.multicast), which halves the FP64-bound generation per output. Measured on C1: Gaussian
    // 1.74 -> 1.31 ms, uniform 0.76 -> 0.71 ms; clusters of 4 lose more to lock-step coupling than they save (1.92 ms).
    // Generating several K blocks per pass for more independent FP64 chains per thread was tried and gained < 5 %.
    a.cl = (a.tiles_n % 2 == 0) ? 2 : 1;
    if (const char* ev = getenv("RBX_TC_CLUSTER")) { const int v = atoi(ev); if ((v == 1 || v == 2 || v == 4) && a.tiles_n % v == 0) a.cl = v; }
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(tiles * ksplit)); cfg.blockDim = dim3(TC_THREADS); cfg.dynamicSmemBytes = TC_SMEM; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned) a.cl; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, kern, tmap, a) != cudaSuccess) { cudaGetLastError(); if (a.partial) cudaFreeAsync(a.partial, stream); return RBX_ERR_CUDA; }
    }
    count_launch();
    int rc = check_launch("sketch_dense_tc_kernel");
    if (rc == RBX_OK && ksplit > 1) {
        int64_t blocks = (g.M * g.N + 255) / 256;
        if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
        tc_reduce_kernel<<<(unsigned) blocks, 256, 0, stream>>>(a.partial, ksplit, g.M, g.N, a.Mpad, a.Npad, g.alpha, g.beta, g.C, g.sc_i, g.sc_j);
        count_launch();
        rc = check_launch("tc_reduce_kernel");
    }
    if (a.partial) cudaFreeAsync(a.partial, stream);
    return rc;
}

// Data with unit stride along j instead of along the contraction axis (e.g. RowMajor A in a left sketch, ColMajor A in a
// right sketch): the tcgen05 kernel wants K-major tiles, so the data is copied chunk by chunk along k into a k-contiguous
// workspace (at most 512 MB, whatever K is) and every chunk takes the normal path with the operator window advanced and
// beta = 1 after the first. The copy is one extra pass over A (0.1 ms per 268 MB) against a 16x slower SIMT fallback.
static int tc_via_kmajor_copy(const GemmOperands<float>& g, cudaStream_t stream) {
    if (g.N > 65535ll * 64) return RBX_ERR_UNSUPPORTED;
    int64_t budget = (int64_t) 512 << 20;
    if (const char* ev = getenv("RBX_TC_COPY_MB")) { const int64_t v = atoll(ev); if (v >= 1) budget = v << 20; }     // tests: several chunks at small sizes
    // Many columns: chunk along j with the whole contraction axis in the workspace, so that every column block of C is written
    // once. (Chunks along k make every chunk after the first read C back and add to it: at 2048 x 4096 x 65536 that doubled the
    // time, at 256 x 4096 x 262144 it was 5.5x.) An implicit Gaussian operator is regenerated per block -- 169 / nb of the block's
    // tensor time -- hence the lower bound on the block width.
    const int64_t kpad = (g.K + 31) / 32 * 32;
    const int64_t nb = (budget / (4 * kpad)) / TC_BN * TC_BN;
    if ((nb >= 1024 || nb >= g.N) && nb >= TC_BN && !getenv("RBX_TC_COPY_ALONG_K")) {
        const int64_t wcols = nb < g.N ? nb : g.N;
        float* ws = nullptr;
        if (ws_alloc(&ws, sizeof(float) * (size_t) wcols * (size_t) kpad, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
        int rc = RBX_OK;
        for (int64_t j0 = 0; j0 < g.N && rc == RBX_OK; j0 += nb) {
            const int64_t nl = (j0 + nb <= g.N) ? nb : g.N - j0;
            rc = launch_kmajor_copy<float>(g.X + j0 * g.sx_j, g.K, nl, g.sx_l, g.sx_j, ws, kpad, stream);
            if (rc != RBX_OK) break;
            GemmOperands<float> gp = g;
            gp.X = ws; gp.sx_l = 1; gp.sx_j = kpad; gp.N = nl; gp.C = g.C + j0 * g.sc_j;
            if (nl < RBX_TC_MIN_N) { rc = launch_sketch_dense_simt<float>(gp, stream); continue; }     // a last block of one column
            rc = launch_sketch_dense_tc_f32(gp, stream);
            if (rc == RBX_ERR_UNSUPPORTED) rc = launch_sketch_dense_simt<float>(gp, stream);          // a narrow last block below the work threshold
        }
        cudaFreeAsync(ws, stream);
        return rc;
    }
    int64_t kc = (budget / (4 * g.N)) / 32 * 32;
    if (kc < 512) return RBX_ERR_UNSUPPORTED;
    const int64_t nch = (g.K + kc - 1) / kc;
    kc = ((g.K + nch - 1) / nch + 31) / 32 * 32;                          // balanced chunks: none is shorter than kc / 2
    float* ws = nullptr;
    if (ws_alloc(&ws, sizeof(float) * (size_t) g.N * (size_t) kc, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    int rc = RBX_OK;
    for (int64_t l0 = 0; l0 < g.K && rc == RBX_OK; l0 += kc) {
        const int64_t kl = (l0 + kc <= g.K) ? kc : g.K - l0;
        rc = launch_kmajor_copy<float>(g.X + l0 * g.sx_l, kl, g.N, g.sx_l, g.sx_j, ws, kc, stream);   // transposing copy, or a plain repack of misaligned k-contiguous data
        if (rc != RBX_OK) break;
        GemmOperands<float> gp = g;
        gp.X = ws; gp.sx_l = 1; gp.sx_j = kc; gp.K = kl;
        gp.beta = (l0 == 0) ? g.beta : 1.f;
        if (g.Sx) gp.Sx = g.Sx + l0 * g.ss_l;
        else if (g.l_on_cols) gp.w.pc0 = g.w.pc0 + l0;
        else gp.w.pr0 = g.w.pr0 + l0;
        rc = launch_sketch_dense_tc_f32(gp, stream);
        if (rc == RBX_ERR_UNSUPPORTED && l0 > 0) rc = RBX_ERR_CUDA;      // cannot happen (same shape as the first chunk); never leave C half-updated silently
    }
    cudaFreeAsync(ws, stream);
    return rc;
}

// DenseSkOp::buff in a form the kernel cannot read directly (unit stride along i: a short-axis-major operator; a row length
// or base that is not 16-byte aligned: an odd operator dimension): copied panel by panel into an aligned k-contiguous
// workspace (<= 512 MB) that the 'X' form of the kernel reads like a generated panel.
static int tc_via_operator_copy(const GemmOperands<float>& g, cudaStream_t stream) {
    if (g.ss_l != 1 && g.ss_i != 1) return RBX_ERR_UNSUPPORTED;
    int64_t kp = (((int64_t) 512 << 20) / (4 * g.M)) / 32 * 32;
    if (kp < 256) kp = 256;
    if (kp > g.K) kp = (g.K + 31) / 32 * 32;
    float* ws = nullptr;
    if (ws_alloc(&ws, sizeof(float) * (size_t) g.M * (size_t) kp, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
    int rc = RBX_OK;
    GemmOperands<float> gp = g;
    for (int64_t l0 = 0; l0 < g.K && rc == RBX_OK; l0 += kp) {
        const int64_t kl = (l0 + kp <= g.K) ? kp : g.K - l0;
        rc = launch_kmajor_copy<float>(g.Sx + l0 * g.ss_l, kl, g.M, g.ss_l, g.ss_i, ws, kp, stream);
        if (rc != RBX_OK) break;
        gp.K = kl; gp.X = g.X + l0; gp.beta = (l0 == 0) ? g.beta : 1.f;
        rc = tc_launch(gp, 'X', ws, kp, stream);
    }
    cudaFreeAsync(ws, stream);
    return rc;
}

int launch_sketch_dense_tc_f32(const GemmOperands<float>& g, cudaStream_t stream) {
    if (getenv("RBX_NO_TC")) return RBX_ERR_UNSUPPORTED;
    // shortest contraction axis worth the kernel. A short contraction ("lift") is all epilogue; the epilogue holds one row of C per
    // TMEM lane: coalesced when C's unit stride is along i, 16-byte vector stores when it is along j and the rows are 16-byte
    // aligned, scalar and uncoalesced otherwise (measured cross-over K ~ 80 there, tools/micro/lift_dense.py)
    const bool c_rows_vec = g.sc_j == 1 && (g.sc_i & 3) == 0 && (reinterpret_cast<uintptr_t>(g.C) & 15) == 0;
    const int64_t min_k = (g.sc_i == 1 || c_rows_vec) ? RBX_TC_MIN_K : 96;
    const bool j_major = g.sx_l != 1 && g.sx_j == 1;
    const bool k_major_misaligned = g.sx_l == 1 && g.N >= 2 && g.N <= 65535 && ((g.sx_j * 4) % 16 != 0 || reinterpret_cast<uintptr_t>(g.X) % 16 != 0);
    if ((j_major || k_major_misaligned) && g.N >= RBX_TC_MIN_N && g.M >= 1 && g.K >= min_k && (double) g.M * g.N * g.K >= dense_min_macs(1e8) &&
        !getenv("RBX_TC_NO_COPY"))
        return tc_via_kmajor_copy(g, stream);
    if (g.sx_l != 1 || g.N < 2 || (g.sx_j * 4) % 16 != 0 || reinterpret_cast<uintptr_t>(g.X) % 16 != 0) return RBX_ERR_UNSUPPORTED;
    if (g.M < 1 || g.N < RBX_TC_MIN_N || g.K < min_k || (double) g.M * g.N * g.K < dense_min_macs(1e8)) return RBX_ERR_UNSUPPORTED;
    if (g.K > 0x7fffffffll - TC_BK || g.N > 0x7fffffffll - TC_BN) return RBX_ERR_UNSUPPORTED;
    if (g.alpha == 0.f) return launch_scale_view<float>(g.C, g.M, g.N, g.sc_i, g.sc_j, g.beta, stream);
    if (g.Sx) {      // DenseSkOp::buff given: k-contiguous rows, 16-byte aligned, no ragged float4 at the end of a row
        if (g.ss_l != 1 || g.ss_i % 4 != 0 || g.K % 4 != 0 || reinterpret_cast<uintptr_t>(g.Sx) % 16 != 0) return tc_via_operator_copy(g, stream);
        return tc_launch(g, 'X', g.Sx, g.ss_i, stream);
    }
    // Implicit operator. The fused kernel generates the operator tile from the counter inside the CTA; it needs the parent
    // rows along k and a 4-aligned column offset. The Gaussian family is bound by the generator (bit-exact Box-Muller,
    // FP64) because every column tile regenerates the tile at 16 warps per SM, so by default it goes through k-panels
    // instead when there are more than two column tiles (a CTA pair shares one generated tile): K1 generates each panel
    // ONCE at full occupancy into a fixed-size float workspace (<= 512 MB however large d * m is; two buffers and an
    // auxiliary stream when there are several panels) and the same tcgen05 kernel reads it. Measured at C1: fused 1.30 ms,
    // one panel 0.96 ms (generate 0.18 + contract 0.75 + reduce 0.02); shorter panels lose more to the per-launch ramp and
    // split-K epilogue than the overlap of generation and contraction returns (4 x 64 MB panels: 1.07 ms).
    const bool fused_ok = g.l_on_cols && (g.w.pc0 & 3) == 0;
    bool use_panel = !fused_ok || (g.w.family == 'G' && (g.N + TC_BN - 1) / TC_BN > 2);
    if (const char* ev = getenv("RBX_TC_PANEL")) { if (atoi(ev) == 0 && fused_ok) use_panel = false; else if (atoi(ev) == 1) use_panel = true; }
    if (!use_panel) return tc_launch(g, g.w.family, nullptr, 0, stream);

    int64_t kp = ((int64_t) 512 << 20) / (4 * g.M);
    if (const char* ev = getenv("RBX_TC_PANEL_K")) { const int64_t v = atoll(ev); if (v >= 32) kp = v; }
    kp = (kp / 32) * 32;
    if (kp < 256) kp = 256;
    if (kp > g.K) kp = (g.K + 31) / 32 * 32;
    const int64_t npanels = (g.K + kp - 1) / kp;
    const int nbuf = npanels > 1 ? 2 : 1;
    const size_t panel_elems = (size_t) g.M * (size_t) kp;
    float* panel = nullptr;
    if (ws_alloc(&panel, sizeof(float) * panel_elems * nbuf, stream) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_ALLOC; }
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
    GemmOperands<float> gp = g;
    int64_t p = 0;
    for (int64_t l0 = 0; l0 < g.K && rc == RBX_OK; l0 += kp, ++p) {
        const int64_t kl = (l0 + kp <= g.K) ? kp : g.K - l0;
        const int b = (nbuf == 2) ? (int)(p & 1) : 0;
        float* buf = panel + (size_t) b * panel_elems;
        if (nbuf == 2 && p >= 2 && cudaStreamWaitEvent(aux, used[b], 0) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; break; }
        if (g.l_on_cols) rc = launch_fill_window<float>(g.w.family, g.w.W * 4, buf, g.M, kl, g.w.pr0, g.w.pc0 + l0, g.w.ctr, g.w.key,
                                                        kp, 1, g.w.family == 'U', aux);
        else             rc = launch_fill_window<float>(g.w.family, g.w.W * 4, buf, kl, g.M, g.w.pr0 + l0, g.w.pc0, g.w.ctr, g.w.key,
                                                        1, kp, g.w.family == 'U', aux);
        if (rc != RBX_OK) break;
        if (nbuf == 2 && (cudaEventRecord(filled[b], aux) != cudaSuccess || cudaStreamWaitEvent(stream, filled[b], 0) != cudaSuccess)) {
            cudaGetLastError(); rc = RBX_ERR_CUDA; break;
        }
        gp.K = kl; gp.X = g.X + l0; gp.beta = (l0 == 0) ? g.beta : 1.f;
        rc = tc_launch(gp, 'X', buf, kp, stream);
        if (rc == RBX_OK && nbuf == 2 && cudaEventRecord(used[b], stream) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    }
    if (nbuf == 2) {
        if (rc != RBX_OK) cudaStreamSynchronize(aux);
        if (ready) cudaEventDestroy(ready);
        for (int b = 0; b < 2; ++b) { if (filled[b]) cudaEventDestroy(filled[b]); if (used[b]) cudaEventDestroy(used[b]); }
    }
    cudaFreeAsync(panel, stream);
    return rc;
}

}  // namespace rbx
