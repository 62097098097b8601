// K4: streaming SASO sketch  B (d x n) = alpha * S (d x m) * A (m x n) + beta * B,  S = wide Short-axis sparse operator
// with vec_nnz signed ones per column (the lskges hot case, BASELINE config 2), A in either layout.
//
// Replaces sparse::lskges -> left_spmm -> coo::apply_coo_left_jki_p11 (skge.hh:485-510, spmm_dispatch.hh:48-160,
// coo_spmm_impl.hh:79-162): the reference sorts the triplets on the host and makes n passes over them.
//
// Design (what was measured on B200 is in DESIGN.md, "SASO kernel"):
//   * a CTA owns a tile of B, 1024 rows x one 128-byte row segment (16 doubles / 32 floats), resident in REGISTERS for
//     the whole pass: warp w owns 64 rows; a lane owns one column of the segment and (double: its half-warp's) 32 / 64
//     rows, one accumulator register each;
//   * A streams through a shared-memory ring in chunks of CK rows (cp.async for row-contiguous data; vector loads and
//     an XOR-swizzled transposing store for column-contiguous data -- the ring layout is the same for both);
//   * accumulator registers cannot be indexed dynamically, so the operator is delivered SORTED BY ROW: a prep kernel
//     regenerates each CK-column chunk of S on the device (bit-exact Fisher-Yates, rbx_fy.cuh), sorts its entries by
//     (row, column) and emits one 16-byte RECORD per accumulator slot: the first three 32-bit codes of the row (ring
//     offset of the A row | sign << 31) and the entry count; the rare entries beyond three go to a small tail region.
//     The main loop is a statically unrolled walk over the slots: one broadcast LDS.128 (record), then per entry one
//     128-byte gather LDS, one sign flip, one FP add. Nothing in the loop depends on a global load. No atomics, no
//     shared-memory read-modify-write, fixed summation order => run-to-run identical results;
//   * double precision: the two half-warps work on different rows, each with its own record; a row with fewer entries
//     than its partner is padded with a code that points at a zero row of the ring;
//   * the row blocks (d / 1024) of one column segment are adjacent CTAs and re-read the A chunk through L2, so HBM
//     sees each byte of A once.
//   * warp roles: 16 consumer warps walk the slots; a producer warp group issues every cp.async of the ring and is coupled
//     to them only through two mbarriers per stage (full / empty), so a warp that finishes its slots early moves on to
//     the next chunk instead of waiting at a CTA-wide barrier for the slowest of the 16 (that barrier cost 20 %);
//     setmaxnreg gives the producers' registers to the consumers.
// Rejected after measurement: predicated "first tail element of every slot in flight together" rounds before the serial
// tail loops (1 / 2 / 3 rounds: +4 / +9 / +20 % time: the kernel is bound by issue slots and shared-memory wavefronts, not
// by the latency of the tail chain), a pause in the producers' polling loop (no change), carrying the gather predicate to
// the add instead of zeroing the value register (ptxas spills ~1 KB per thread),
// TMA boxes for the A chunk (256-byte box rows are TMA row-rate bound, ~28 cycles/row),
// cluster multicast (73 cycles/row), register tiles selected with brx.idx (~45 issue slots per select), B tile in
// shared memory with read-modify-write (latency bound at 16 warps/SM, 24 B of shared-memory traffic per update).
#include <cmath>
#include <cstdlib>
#include "rbx_kernels.cuh"
#include "rbx_fy.cuh"
#include "rbx_tma.cuh"

namespace rbx {

constexpr int SASO_ROWB = 128;      // bytes of one row segment of A / B handled by a CTA
constexpr int SASO_WARPS = 16;
constexpr int SASO_RW = 64;         // rows of B per warp
constexpr int SASO_R = SASO_WARPS * SASO_RW;   // rows of B per CTA
constexpr int SASO_THREADS = SASO_WARPS * 32;
constexpr int SASO_MAX_RB = 8;      // row blocks of 1024 rows: operator windows of up to 8192 rows take the streaming kernel
// ring stages: row-contiguous data runs two long chunks (compute per chunk far exceeds the copy latency, and a longer
// chunk means more entries per accumulator slot); the slower element-wise copy of column-contiguous data wants three
#ifndef SASO_NST_COLS_VAL
#define SASO_NST_COLS_VAL 2
#endif
#ifndef SASO_CK_MAX
#define SASO_CK_MAX 704         // longest chunk (a multiple of 64); the ring stages must fit in shared memory
#endif
#ifndef SASO_CK_MAX_COLS
#define SASO_CK_MAX_COLS 576    // column-contiguous data (measured at C2: 704 -> 11.1 ms, 576 -> 10.8 ms, 512 -> 11.0 ms)
#endif
#ifndef SASO_ALWAYS
#define SASO_ALWAYS 1         // leading gathers of a slot that are not predicated (a row has a first entry 86 % of the time at C2)
#endif
constexpr int SASO_NST_ROWS = 2, SASO_NST_COLS = SASO_NST_COLS_VAL;
constexpr uint32_t SASO_SIGN = 0x80000000u;
#ifndef SASO_NU_VAL
#define SASO_NU_VAL 3
#endif
constexpr int SASO_NU = SASO_NU_VAL;          // codes held by a slot record; the rest of a slot's elements are its tail
#ifndef SASO_PRED_ADD
#define SASO_PRED_ADD 0   // 1: padded gathers carry their predicate to the add instead of zeroing the value register
#endif
#ifndef SASO_POLL_NS
#define SASO_POLL_NS 0    // > 0: the producer warps sleep this many ns between polls of a stage's empty barrier
#endif
#ifndef SASO_GROUP
#define SASO_GROUP 4      // slots whose records and gathers are in flight together
#endif
static_assert(SASO_NU == 3, "the 16-byte slot record holds three codes and the count");
// SASO_PRODUCER = 1: a 17th warp issues every cp.async of the ring and the 16 consumer warps are coupled to it only through
// two mbarriers per stage (full: the copies have landed; empty: all consumer warps are done with the stage), so a warp that
// finishes its slots early starts on the next chunk instead of waiting at a CTA-wide barrier for the slowest warp.
#ifndef SASO_PRODUCER
#define SASO_PRODUCER 2
#endif
// SASO_PRODUCER = 2: a whole producer warp group (registers are allocated to groups of four warps anyway), which hands most
// of its registers to the consumers with setmaxnreg.
// The CTA's register pool is what it was launched with: 640 threads x 96. 16 consumer warps x 112 + 4 producer warps x 24
// fits; 120 does not and the allocation would spin forever.
#ifndef SASO_CONS_REGS
#define SASO_CONS_REGS 112
#endif
static_assert(SASO_PRODUCER != 2 || SASO_WARPS * SASO_CONS_REGS + 4 * 24 <= (SASO_WARPS + 4) * 96, "setmaxnreg.inc beyond the CTA's register pool deadlocks");
#ifndef SASO_BULK_META
#define SASO_BULK_META 1          // 1: with row-contiguous data the slot records and tail region of a chunk arrive by two bulk copies (TMA)
                                  // instead of 16-byte cp.async (C2: 9.40 -> 9.23 ms; column-contiguous data lost 3 % with it and keeps cp.async)
#endif
constexpr int SASO_PROD_WARPS = SASO_PRODUCER == 2 ? 4 : (SASO_PRODUCER ? 1 : 0);
constexpr int SASO_TILE_THREADS = SASO_THREADS + SASO_PROD_WARPS * 32;
constexpr int SASO_LOAD_THREADS = SASO_PRODUCER ? SASO_PROD_WARPS * 32 : SASO_THREADS;    // threads that share the copies of one chunk

// exclusive scan of in[0..n) into out[0..n], out[n] = total; all SASO_THREADS threads of the block call it
__device__ __forceinline__ void block_excl_scan(const uint32_t* in, uint32_t* out, int n, uint32_t* wtot) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (n + SASO_THREADS - 1) / SASO_THREADS;
    const int b = t * per, e = (b + per < n) ? b + per : n;
    uint32_t local = 0;
    for (int i = b; i < e; ++i) local += in[i];
    uint32_t inc = local;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += y; }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < SASO_WARPS ? wtot[lane] : 0, winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= d) winc += y; }
        if (lane < SASO_WARPS) wtot[lane] = winc - w;
        if (lane == SASO_WARPS - 1) out[n] = winc;
    }
    __syncthreads();
    uint32_t run = inc - local + wtot[warp];
    for (int i = b; i < e; ++i) { const uint32_t v = in[i]; out[i] = run; run += v; }
    __syncthreads();
}

// ---- prep: one block per chunk of CK columns of S -----------------------------------------------------------
// Slot q = bucket * NS + s owns rows bucket*64 + h*NS + s, h < HALVES (the half-warps of a double-precision warp work on
// different rows). n(q) = max over h of the row's entry count in this chunk. Outputs, per chunk:
//   rec[chunk * rec_stride + (q * HALVES + h) * 4 ..]   one 16-byte record per (slot, half): the row's first SASO_NU
//        codes (zero code where it has fewer) and the slot's tail count max(n(q) - SASO_NU, 0). The record of a bucket's
//        slot 0 also carries, above bit 12, where the bucket's tail codes start inside its row block's region, and in bit 31 "not in the region";
//   stream[chunk * stream_stride + (rb * cap + ...) * HALVES + h]   region of row block rb: the tails (elements SASO_NU ..
//        n(q)-1, HALVES codes per element, the shorter row padded with the zero code) of its buckets, bucket after bucket,
//        as long as they fit in cap elements. The main kernel copies the region into the ring stage;
//   stream[chunk * stream_stride + (nrb * cap + ws[b] + ...) * HALVES + h]   the tails of buckets that did not fit (far
//        tail of the distribution), read straight from global memory; ws[chunk * ws_stride + b] = their start.
template <int KMAX, int HALVES>
__global__ void __launch_bounds__(SASO_THREADS) saso_prep_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t ro, int64_t co,
                                                               int64_t M, int64_t K, int CK, int nbuckets, int nrb, int cols_contig, int cap, int R,
                                                               uint32_t* __restrict__ stream, int stream_stride,
                                                               uint32_t* __restrict__ rec, int rec_stride,
                                                               uint32_t* __restrict__ ws, int ws_stride, int* __restrict__ spill_flag) {
    constexpr int NS = SASO_RW / HALVES;
    extern __shared__ uint32_t sh32[];
    __shared__ uint32_t wtot[SASO_WARPS];
    const int t = threadIdx.x;
    const int Mp = nbuckets * SASO_RW, E = CK * vec_nnz, nslots = nbuckets * NS;
    uint32_t* sorted = sh32;                 // E      codes ordered by (row, column)
    uint32_t* tcode = sorted + E;            // E      codes in generation order
    uint32_t* rowoff = tcode + E;            // Mp + 1
    uint32_t* rowcnt = rowoff + Mp + 1;      // Mp
    uint32_t* slotcnt = rowcnt + Mp;         // nslots
    uint32_t* slotoff = slotcnt + nslots;    // nslots + 1
    uint32_t* tailcnt = slotoff + nslots + 1;   // nslots
    uint16_t* trow = reinterpret_cast<uint16_t*>(tailcnt + nslots);       // E
    const int64_t chunk = blockIdx.x;
    // a code is the ring offset of A row c in elements: c * (row pitch / sizeof(T)); the pitch is 128 bytes, plus one
    // element when the data is column-contiguous (see saso_tile_kernel). Row CK of the ring is all zeros.
    const uint32_t pitch_el = 32u / HALVES + (cols_contig ? 1u : 0u);
    const uint32_t zero_code = (uint32_t) CK * pitch_el;

    uint32_t* so = stream + chunk * stream_stride;
    for (int r = t; r < Mp; r += SASO_THREADS) rowcnt[r] = 0;
    for (int e = t; e < E; e += SASO_THREADS) trow[e] = 0xFFFF;
    __syncthreads();
    for (int cc = t; cc < CK; cc += SASO_THREADS) {
        const int64_t l = chunk * CK + cc;
        if (l >= K) break;
        const uint32_t base = (uint32_t) cc * pitch_el;
        fy_vector<KMAX>(ctr, key, co + l, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
            int64_t r = idx - ro;
            if (r >= 0 && r < M) {
                r += (int64_t)(cc % R) * M;      // R > 1: column cc of the chunk feeds copy (cc mod R) of the M output rows (see launch_saso_fast)
                tcode[cc * vec_nnz + j] = base | (negative ? SASO_SIGN : 0u);
                trow[cc * vec_nnz + j] = (uint16_t) r;
                atomicAdd(&rowcnt[r], 1u);
            }
        });
    }
    __syncthreads();
    block_excl_scan(rowcnt, rowoff, Mp, wtot);
    for (int r = t; r < Mp; r += SASO_THREADS) rowcnt[r] = 0;
    __syncthreads();
    for (int e = t; e < E; e += SASO_THREADS) {
        const uint32_t r = trow[e];
        if (r != 0xFFFF) sorted[rowoff[r] + atomicAdd(&rowcnt[r], 1u)] = tcode[e];
    }
    __syncthreads();
    // fixed order inside a row (by column): the scatter above is in arrival order
    for (int r = t; r < Mp; r += SASO_THREADS) {
        uint32_t* s = sorted + rowoff[r];
        const int n = (int) rowcnt[r];
        for (int i = 1; i < n; ++i) {
            const uint32_t v = s[i];
            int k = i - 1;
            while (k >= 0 && (s[k] & ~SASO_SIGN) > (v & ~SASO_SIGN)) { s[k + 1] = s[k]; --k; }
            s[k + 1] = v;
        }
    }
    for (int q = t; q < nslots; q += SASO_THREADS) {
        const int b = q / NS, s = q % NS;
        uint32_t n = 0;
#pragma unroll
        for (int h = 0; h < HALVES; ++h) { const uint32_t v = rowcnt[b * SASO_RW + h * NS + s]; n = v > n ? v : n; }
        slotcnt[q] = n;
        tailcnt[q] = n > (uint32_t) SASO_NU ? n - (uint32_t) SASO_NU : 0u;
    }
    __syncthreads();
    block_excl_scan(tailcnt, slotoff, nslots, wtot);
    // first tail element of each row block, and whether any bucket of the chunk spills out of its region
    uint32_t rb_base[SASO_MAX_RB];
    bool any_spill = false;
    for (int rb = 0; rb < SASO_MAX_RB; ++rb) {
        const int q0 = rb * SASO_WARPS * NS, q1 = (rb + 1) * SASO_WARPS * NS;
        rb_base[rb] = slotoff[q0 < nslots ? q0 : nslots];
        any_spill = any_spill || (slotoff[q1 < nslots ? q1 : nslots] - rb_base[rb] > (uint32_t) cap);
    }
    const uint32_t region_words = (uint32_t)(nrb * cap * HALVES);
    for (uint32_t e = t; e < region_words; e += SASO_THREADS) so[e] = zero_code;
    if (any_spill) {
        if (t == 0) *spill_flag = 1;
        const uint32_t used = slotoff[nslots] * HALVES;
        for (uint32_t e = t; e < used; e += SASO_THREADS) so[region_words + e] = zero_code;
    }
    uint32_t* wo = ws + chunk * ws_stride;
    for (int b = t; b <= nbuckets; b += SASO_THREADS) wo[b] = slotoff[b * NS];
    __syncthreads();
    uint32_t* ro_ = rec + chunk * rec_stride;
    for (int q = t; q < rec_stride / (4 * HALVES); q += SASO_THREADS) {
        const bool live = q < nslots;
        const int b = q / NS, s = q % NS, rb = b / SASO_WARPS;
        uint32_t n = 0, w = 0;
        uint32_t* tails = so;
        if (live) {
            n = slotcnt[q];
            const uint32_t b_start = slotoff[b * NS] - rb_base[rb], b_end = slotoff[(b + 1) * NS] - rb_base[rb];
            const bool fits = b_end <= (uint32_t) cap;
            tails = fits ? so + ((size_t) rb * cap + (slotoff[q] - rb_base[rb])) * HALVES : so + region_words + (size_t) slotoff[q] * HALVES;
            w = tailcnt[q] | (s == 0 ? ((fits ? b_start << 12 : SASO_SIGN)) : 0u);
        }
#pragma unroll
        for (int h = 0; h < HALVES; ++h) {
            const int row = b * SASO_RW + h * NS + s;
            const uint32_t nr = live ? rowcnt[row] : 0u;
            const uint32_t* src = sorted + (live ? rowoff[row] : 0u);
            uint32_t r4[3];
#pragma unroll
            for (int i = 0; i < SASO_NU; ++i) r4[i] = (uint32_t) i < nr ? src[i] : zero_code;
            *reinterpret_cast<uint4*>(ro_ + ((size_t) q * HALVES + h) * 4) = make_uint4(r4[0], r4[1], r4[2], w);
            for (uint32_t i = SASO_NU; i < nr; ++i) tails[(i - SASO_NU) * HALVES + h] = src[i];
        }
    }
}

template <typename T>
struct SasoArgs {
    const T* X; int64_t sx_l, sx_j; int64_t K, N;      // data view: X[l * sx_l + j * sx_j], one of the strides is 1
    const uint32_t* stream; const uint32_t* rec; const uint32_t* ws;
    int stream_stride, rec_stride, ws_stride;
    int CK, cap;                                       // rows per chunk; tail elements a row block's region holds
    int nbuckets, njb, nrb, msplit, nchunks, dbg_skip;
    int x_elem;                                        // row-contiguous data whose rows are not 16-byte aligned: element-wise copies
    int direct;                                        // 1: one split and no row copies -> the tile kernel stores alpha * acc + beta * C itself
    int64_t M;                                         // rows the tile kernel owns (R copies of the window's rows when R > 1)
    T alpha, beta;
    T* C; int64_t sc_i, sc_j;
    T* partial;                                        // [msplit][M][njb * W] when msplit > 1
    const int* spill_flag;                             // set by the prep kernel: some tail codes are outside the regions
};

// asynchronous global->shared copies; bytes beyond src_bytes are zero-filled (ragged edges)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
template <int BYTES> __device__ __forceinline__ void cp_async_small(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" :: "r"(dst), "l"(src), "n"(BYTES), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// (+/-) *(T*)addr from shared memory, the sign taken from bit 31 of code
template <typename T> __device__ __forceinline__ T lds_signed(uint32_t addr, uint32_t code);
template <> __device__ __forceinline__ double lds_signed<double>(uint32_t addr, uint32_t code) {
    double v;
    asm volatile("{\n\t.reg .b32 lo, hi;\n\tld.shared.v2.b32 {lo, hi}, [%1];\n\tlop3.b32 hi, hi, %2, 0x80000000, 0x78;\n\tmov.b64 %0, {lo, hi};\n\t}"
                 : "=d"(v) : "r"(addr), "r"(code));
    return v;
}
template <> __device__ __forceinline__ float lds_signed<float>(uint32_t addr, uint32_t code) {
    float v;
    asm volatile("{\n\t.reg .b32 x;\n\tld.shared.b32 x, [%1];\n\tlop3.b32 x, x, %2, 0x80000000, 0x78;\n\tmov.b32 %0, x;\n\t}"
                 : "=f"(v) : "r"(addr), "r"(code));
    return v;
}
// raw bits of *(T*)addr from shared memory, or zero (and no shared-memory access by this lane) when code is the zero code
template <typename T> struct RawBits;
template <> struct RawBits<double> { typedef unsigned long long type; };
template <> struct RawBits<float> { typedef uint32_t type; };
template <typename T> __device__ __forceinline__ typename RawBits<T>::type lds_raw_nz(uint32_t addr, uint32_t code, uint32_t zero_code);
template <> __device__ __forceinline__ unsigned long long lds_raw_nz<double>(uint32_t addr, uint32_t code, uint32_t zero_code) {
    unsigned long long v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, %3;\n\tmov.b64 %0, 0;\n\t@p ld.shared.b64 %0, [%1];\n\t}" : "=l"(v) : "r"(addr), "r"(code), "r"(zero_code));
    return v;
}
template <> __device__ __forceinline__ uint32_t lds_raw_nz<float>(uint32_t addr, uint32_t code, uint32_t zero_code) {
    uint32_t v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, %3;\n\tmov.b32 %0, 0;\n\t@p ld.shared.b32 %0, [%1];\n\t}" : "=r"(v) : "r"(addr), "r"(code), "r"(zero_code));
    return v;
}
// acc += (+/-) *(T*)addr unless code is the zero code: predicate, load, sign flip and add in one predicated sequence (no
// zeroing of the value register; the predicate lives from the compare to the add). SASO_PRED_ADD selects it.
template <typename T> __device__ __forceinline__ void lds_add_nz(T& acc, uint32_t addr, uint32_t code, uint32_t zero_code);
template <> __device__ __forceinline__ void lds_add_nz<double>(double& acc, uint32_t addr, uint32_t code, uint32_t zero_code) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 lo, hi;\n\t.reg .f64 v;\n\tsetp.ne.u32 p, %2, %3;\n\t@p ld.shared.v2.b32 {lo, hi}, [%1];\n\t"
                 "@p lop3.b32 hi, hi, %2, 0x80000000, 0x78;\n\t@p mov.b64 v, {lo, hi};\n\t@p add.f64 %0, %0, v;\n\t}"
                 : "+d"(acc) : "r"(addr), "r"(code), "r"(zero_code));
}
template <> __device__ __forceinline__ void lds_add_nz<float>(float& acc, uint32_t addr, uint32_t code, uint32_t zero_code) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 x;\n\t.reg .f32 v;\n\tsetp.ne.u32 p, %2, %3;\n\t@p ld.shared.b32 x, [%1];\n\t"
                 "@p lop3.b32 x, x, %2, 0x80000000, 0x78;\n\t@p mov.b32 v, x;\n\t@p add.f32 %0, %0, v;\n\t}"
                 : "+f"(acc) : "r"(addr), "r"(code), "r"(zero_code));
}
__device__ __forceinline__ double apply_sign(unsigned long long bits, uint32_t code) {
    return __hiloint2double((int)((uint32_t)(bits >> 32) ^ (code & SASO_SIGN)), (int)(uint32_t) bits);
}
__device__ __forceinline__ float apply_sign(uint32_t bits, uint32_t code) { return __uint_as_float(bits ^ (code & SASO_SIGN)); }

// slots [0, NS) of this warp: acc[s] += sum of the gathered (signed) A values of its elements.
// recs: shared address of this lane's first 16-byte slot record; region_s: shared address of this lane's first code of the
// row block's tail region in the stage; spill / wsp: where the tails are in global memory when the bucket's did not fit in
// the region (SPILL variant only); lanebase: shared address of this lane's column in row 0 of the stage's A chunk.
// Address of an element = lanebase + (code << log2(sizeof(T))); the shift drops the sign bit.
// The walk goes G slots at a time: their records, then all of their gathers, are in flight together (4 warps per
// scheduler cannot hide a record -> gather -> add chain per slot); the tails follow, one short loop per slot.
template <typename T, int NS, int HALVES, bool SPILL>
__device__ __forceinline__ void saso_consume(T (&acc)[NS], uint32_t recs, uint32_t region_s, const uint32_t* spill, const uint32_t* wsp,
                                             uint32_t lanebase, uint32_t zero_code) {
    constexpr int SH = sizeof(T) == 8 ? 3 : 2;
    constexpr int G = SASO_GROUP;
    typedef typename RawBits<T>::type raw_t;
    const uint32_t* tp = spill;      // SPILL: generic pointer to the bucket's tail codes (the stage's region or global memory)
    uint32_t tps = region_s;         // !SPILL: their shared-memory address
#pragma unroll
    for (int g = 0; g < NS; g += G) {
        uint32_t c[G][4];
#pragma unroll
        for (int j = 0; j < G; ++j)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(c[j][0]), "=r"(c[j][1]), "=r"(c[j][2]), "=r"(c[j][3])
                         : "r"(recs + (uint32_t)((g + j) * HALVES * 16)));
        if (g == 0) {
            const uint32_t w = c[0][3];
            c[0][3] = w & 0xfffu;
            const uint32_t off = ((w >> 12) & 0x7ffffu) * (uint32_t)(HALVES * 4);
            if (SPILL) tp = (w & SASO_SIGN) ? spill + (size_t) __ldg(wsp) * HALVES
                                            : reinterpret_cast<const uint32_t*>(__cvta_shared_to_generic(region_s + off));
            else tps = region_s + off;
        }
        raw_t v[G][SASO_NU];
#pragma unroll
        for (int k = 0; k < SASO_NU; ++k)
#pragma unroll
            for (int j = 0; j < G; ++j) {
#if SASO_PRED_ADD
                if (k >= SASO_ALWAYS) { lds_add_nz<T>(acc[g + j], lanebase + (c[j][k] << SH), c[j][k], zero_code); continue; }
#endif
                if (k < SASO_ALWAYS) {
                    if (sizeof(T) == 8) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(*reinterpret_cast<unsigned long long*>(&v[j][k])) : "r"(lanebase + (c[j][k] << SH)));
                    else asm volatile("ld.shared.b32 %0, [%1];" : "=r"(*reinterpret_cast<uint32_t*>(&v[j][k])) : "r"(lanebase + (c[j][k] << SH)));
                } else {
                    v[j][k] = lds_raw_nz<T>(lanebase + (c[j][k] << SH), c[j][k], zero_code);
                }
            }
#pragma unroll
        for (int k = 0; k < (SASO_PRED_ADD ? SASO_ALWAYS : SASO_NU); ++k)
#pragma unroll
            for (int j = 0; j < G; ++j) acc[g + j] += apply_sign(v[j][k], c[j][k]);
        uint32_t any = 0;
#pragma unroll
        for (int j = 0; j < G; ++j) any |= c[j][3];
        if (any != 0) {                             // some slot of the group has more than SASO_NU elements
#pragma unroll
            for (int j = 0; j < G; ++j) {
                if (SPILL) {
#pragma unroll 1
                    for (uint32_t i = 0; i < c[j][3]; ++i) {
                        uint32_t code;
                        asm volatile("ld.b32 %0, [%1];" : "=r"(code) : "l"(tp));
                        tp += HALVES;
                        acc[g + j] += lds_signed<T>(lanebase + (code << SH), code);
                    }
                } else {
                    const uint32_t end = tps + c[j][3] * (uint32_t)(HALVES * 4);
#pragma unroll 1
                    for (; tps != end; tps += HALVES * 4) {
                        uint32_t code;
                        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(code) : "r"(tps));
                        raw_t raw;
                        if (sizeof(T) == 8) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(*reinterpret_cast<unsigned long long*>(&raw)) : "r"(lanebase + (code << SH)));
                        else asm volatile("ld.shared.b32 %0, [%1];" : "=r"(*reinterpret_cast<uint32_t*>(&raw)) : "r"(lanebase + (code << SH)));
                        acc[g + j] += apply_sign(raw, code);
                    }
                }
            }
        }
    }
}

// COLS_CONTIG: columns of the data view are contiguous (sx_l == 1); otherwise rows are (sx_j == 1).
// Ring layout of one A chunk: row c is 128 bytes at offset c * pitch. Row-contiguous data is copied as is (16-byte
// cp.async, pitch 128); column-contiguous data is copied element by element (8 / 4-byte cp.async, lanes run down a
// column) with a pitch of 128 bytes + one element, which spreads that transposing write over all banks while the gather
// (lanes along a row) stays contiguous.
// SPILL: the variant that can read tail codes from global memory. The prep kernel raises *a.spill_flag when any bucket's
// tails did not fit in its region (far tail of the distribution); both variants are launched and the one the flag does
// not select returns at once. Keeping the global-memory path out of the common variant keeps its unrolled walk small:
// unexecuted code inside that loop cost 2.5 % of the kernel time (instruction cache).
template <typename T, bool COLS_CONTIG, bool SPILL>
__global__ void __launch_bounds__(SASO_TILE_THREADS, 1) saso_tile_kernel(const SasoArgs<T> a) {
    if ((__ldg(a.spill_flag) != 0) != SPILL) return;
    constexpr int W = SASO_ROWB / (int) sizeof(T);     // columns per CTA: 16 doubles / 32 floats
    constexpr int HALVES = 32 / W;                     // entries per warp instruction: 2 / 1
    constexpr int NS = SASO_RW / HALVES;               // accumulators per lane: 32 doubles / 64 floats
    constexpr int EP = 16 / (int) sizeof(T);           // elements per 16-byte piece
    constexpr int REC_BYTES = SASO_WARPS * NS * HALVES * 16;   // slot records of this CTA's 1024 rows, per chunk
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t smem0 = smem_u32(smem);
    const int CK = a.CK;
    constexpr int PITCH = SASO_ROWB + (COLS_CONTIG ? (int) sizeof(T) : 0);
    const int a_bytes = CK * PITCH;
    const int az_bytes = (a_bytes + PITCH + 15) & ~15;                               // A chunk + its zero row
    const int tail_bytes = a.cap * HALVES * 4;
    const int stage_bytes = az_bytes + REC_BYTES + tail_bytes;                       // A | zero row | slot records | tail region

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sub = lane / W, jl = lane % W;
    // the row blocks of one (column segment, split) are adjacent block ids: they run together and share A through L2
    const int rb = blockIdx.x % a.nrb;
    const int cid = blockIdx.x / a.nrb;
    // Row-contiguous data: neighbouring CTAs take neighbouring column segments of the same rows, and a split owns a
    // contiguous range of chunks. Column-contiguous data: every column is its own set of pages, so the splits of one
    // column segment run side by side on interleaved chunks and walk the same pages together (TLB reach).
    const int jb = COLS_CONTIG ? cid / a.msplit : cid % a.njb;
    const int ms = COLS_CONTIG ? cid % a.msplit : cid / a.njb;
    const int per = (a.nchunks + a.msplit - 1) / a.msplit;
    const int cstep = COLS_CONTIG ? a.msplit : 1;
    const int cb = COLS_CONTIG ? ms : ms * per;
    const int ce = COLS_CONTIG ? a.nchunks : ((cb + per < a.nchunks) ? cb + per : a.nchunks);
    const int64_t j0 = (int64_t) jb * W;
    const int bucket = rb * SASO_WARPS + warp;         // warp owns rows [bucket*64, bucket*64+64) of the window
    const bool active = warp < SASO_WARPS && bucket < a.nbuckets;

    constexpr int NST = COLS_CONTIG ? SASO_NST_COLS : SASO_NST_ROWS;
    if (tid < NST * (SASO_ROWB / 8))                   // the zero rows of the stages
        *reinterpret_cast<uint2*>(smem + (size_t)(tid / (SASO_ROWB / 8)) * stage_bytes + a_bytes + (tid % (SASO_ROWB / 8)) * 8) = make_uint2(0, 0);

    // A chunk, slot records and tail region of chunk ch -> stage s. Nothing here depends on data read from global
    // memory: the copies of the next chunk are issued without waiting for anything.
    // Tiles that touch neither edge of A take the copy loops without per-piece bounds arithmetic.
    const bool cols_inside = j0 + W <= a.N;
    constexpr int LT = SASO_LOAD_THREADS;
    const int lt = SASO_PRODUCER ? tid - SASO_THREADS : tid;
    constexpr bool BULK_META = SASO_PRODUCER && SASO_BULK_META && !COLS_CONTIG;
#if SASO_PRODUCER
    __shared__ __align__(8) uint64_t full_bar[NST], empty_bar[NST];
#endif
    auto load_async = [&](int ch, int s) {
        const uint32_t st = smem0 + (uint32_t) s * stage_bytes;
        const bool inside = cols_inside && (int64_t)(ch + 1) * CK <= a.K;
        if (!COLS_CONTIG) {
            if (a.x_elem) {
                // rows that do not start on 16 bytes (an odd leading dimension or base): one element per copy, a warp instruction
                // still moves whole 128-byte row segments
                if (inside) {
                    const T* src = a.X + ((int64_t) ch * CK + lt / W) * a.sx_l + j0 + lt % W;
                    const int64_t step = (int64_t)(LT / W) * a.sx_l;
                    uint32_t dst = st + (lt / W) * SASO_ROWB + (lt % W) * (int) sizeof(T);
#pragma unroll 8
                    for (int i = 0; i < CK / (LT / W); ++i) { cp_async_small<(int) sizeof(T)>(dst, src, (uint32_t) sizeof(T)); dst += (LT / W) * SASO_ROWB; src += step; }
                } else
                for (int p = lt; p < CK * W; p += LT) {
                    const int row = p / W, q = p % W;
                    const int64_t l = (int64_t) ch * CK + row, col = j0 + q;
                    const uint32_t nb = (l < a.K && col < a.N) ? (uint32_t) sizeof(T) : 0u;
                    cp_async_small<(int) sizeof(T)>(st + row * SASO_ROWB + q * (int) sizeof(T), nb ? a.X + l * a.sx_l + col : a.X, nb);
                }
            } else if (inside) {
                const T* src = a.X + ((int64_t) ch * CK + (lt >> 3)) * a.sx_l + j0 + (lt & 7) * EP;
                const int64_t step = (int64_t)(LT / 8) * a.sx_l;
                uint32_t dst = st + lt * 16;
#pragma unroll 8
                for (int i = 0; i < CK / (LT / 8); ++i) { cp_async16(dst, src, 16); dst += LT * 16; src += step; }
            } else {
                for (int p = lt; p < CK * 8; p += LT) {
                    const int row = p >> 3, q = p & 7;
                    const int64_t l = (int64_t) ch * CK + row, col = j0 + (int64_t) q * EP;
                    int64_t nb = (l < a.K) ? (a.N - col) * (int64_t) sizeof(T) : 0;
                    nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
                    const T* src = nb > 0 ? a.X + l * a.sx_l + col : a.X;
                    cp_async16(st + p * 16, src, (uint32_t) nb);
                }
            }
        } else {
            for (int rr = lt; rr < CK; rr += LT) {
                const int64_t l = (int64_t) ch * CK + rr;
                const uint32_t row = st + rr * PITCH;
                if (inside) {
                    const T* src = a.X + l + j0 * a.sx_j;
#pragma unroll
                    for (int jj = 0; jj < W; ++jj) { cp_async_small<(int) sizeof(T)>(row + jj * (int) sizeof(T), src, (uint32_t) sizeof(T)); src += a.sx_j; }
                } else {
                    const T* src = a.X + (l < a.K ? l : 0) + j0 * a.sx_j;
#pragma unroll 4
                    for (int jj = 0; jj < W; ++jj) {
                        const uint32_t nb = (l < a.K && j0 + jj < a.N) ? (uint32_t) sizeof(T) : 0u;
                        cp_async_small<(int) sizeof(T)>(row + jj * (int) sizeof(T), nb ? src + jj * a.sx_j : a.X, nb);
                    }
                }
            }
        }
        const uint8_t* rsrc = reinterpret_cast<const uint8_t*>(a.rec + (size_t) ch * a.rec_stride) + (size_t) rb * REC_BYTES;
        const uint8_t* tsrc = reinterpret_cast<const uint8_t*>(a.stream + (size_t) ch * a.stream_stride) + (size_t) rb * tail_bytes;
#if SASO_PRODUCER
        if (BULK_META) {
          if (lt == 0) {
            mbar_expect_tx(&full_bar[s], (uint32_t)(REC_BYTES + tail_bytes));
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(st + az_bytes), "l"(rsrc), "r"((uint32_t) REC_BYTES), "r"(smem_u32(&full_bar[s])) : "memory");
            if (tail_bytes > 0)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(st + az_bytes + REC_BYTES), "l"(tsrc), "r"((uint32_t) tail_bytes), "r"(smem_u32(&full_bar[s])) : "memory");
          }
        } else
#endif
        {
            for (int p = lt; p < REC_BYTES / 16; p += LT) cp_async16(st + az_bytes + p * 16, rsrc + p * 16, 16);
            // the threads that were idle (column-contiguous) or last (row-contiguous) in the A copy take the tail region
            for (int p = LT - 1 - lt; p < tail_bytes / 16; p += LT) cp_async16(st + az_bytes + REC_BYTES + p * 16, tsrc + p * 16, 16);
        }
    };

    T acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) acc[s] = (T) 0;
    const uint32_t lx = (uint32_t) jl * (uint32_t) sizeof(T);
    const uint32_t zero_code = (uint32_t) CK * (uint32_t)(PITCH / (int) sizeof(T));
    const uint32_t* spill = a.stream + (size_t) a.nrb * a.cap * HALVES + sub;

#if SASO_PRODUCER
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(&full_bar[s], SASO_LOAD_THREADS + (BULK_META ? 1 : 0)); mbar_init(&empty_bar[s], SASO_WARPS); }
        fence_mbar_init();
    }
    __syncthreads();                                    // barriers and zero rows are in place
    if (warp >= SASO_WARPS) {
        // ---- producer warp(s): every copy of the ring ----
#if SASO_PRODUCER == 2
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
#endif
        for (int ch = cb, it = 0; ch < ce; ch += cstep, ++it) {
            const int s = it % NST, u = it / NST;
            if (u > 0) {
#if SASO_POLL_NS > 0
                while (!mbar_try_wait(&empty_bar[s], (uint32_t)(u - 1) & 1u)) __nanosleep(SASO_POLL_NS);
#else
                mbar_wait(&empty_bar[s], (uint32_t)(u - 1) & 1u);
#endif
            }
            load_async(ch, s);
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(smem_u32(&full_bar[s])) : "memory");
        }
        cp_async_commit();
        cp_async_wait<0>();
        return;
    }
#if SASO_PRODUCER == 2
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" :: "n"(SASO_CONS_REGS));
#endif
    for (int ch = cb, it = 0; ch < ce; ch += cstep, ++it) {
        const int s = it % NST;
        mbar_wait(&full_bar[s], (uint32_t)(it / NST) & 1u);
        if (active && a.dbg_skip != 1) {
            const uint32_t st = smem0 + (uint32_t) s * stage_bytes;
            const uint32_t recs = st + az_bytes + (uint32_t)((warp * NS * HALVES + sub) * 16);
            const uint32_t region = st + (uint32_t)(az_bytes + REC_BYTES + sub * 4);
            saso_consume<T, NS, HALVES, SPILL>(acc, recs, region, spill + (size_t) ch * a.stream_stride, a.ws + (size_t) ch * a.ws_stride + bucket,
                                        st + lx, zero_code);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
#else
    constexpr int DIST = NST - 1;                      // prefetch distance in chunks
    for (int s = 0; s < DIST; ++s) {
        if (cb + s * cstep < ce) load_async(cb + s * cstep, s);
        cp_async_commit();
    }
    for (int ch = cb, it = 0; ch < ce; ch += cstep, ++it) {
        cp_async_wait<DIST - 1>();
        __syncthreads();                                // chunk `it` has landed; everyone is done with chunk it-1
        const int nxt = ch + DIST * cstep;
        if (nxt < ce) load_async(nxt, (it + DIST) % NST);
        cp_async_commit();
        if (active && a.dbg_skip != 1) {
            const uint32_t st = smem0 + (uint32_t)(it % NST) * stage_bytes;
            const uint32_t recs = st + az_bytes + (uint32_t)((warp * NS * HALVES + sub) * 16);
            const uint32_t region = st + (uint32_t)(az_bytes + REC_BYTES + sub * 4);
            saso_consume<T, NS, HALVES, SPILL>(acc, recs, region, spill + (size_t) ch * a.stream_stride, a.ws + (size_t) ch * a.ws_stride + bucket,
                                        st + lx, zero_code);
        }
    }
    cp_async_wait<0>();
#endif
    if (active) {
        const int64_t col = j0 + jl;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            const int64_t row = (int64_t) bucket * SASO_RW + sub * NS + s;
            if (row < a.M && col < a.N) {
                if (a.direct) {
                    T* c = a.C + row * a.sc_i + col * a.sc_j;
                    *c = (a.beta == (T) 0) ? a.alpha * acc[s] : a.alpha * acc[s] + a.beta * (*c);
                } else {
                    a.partial[((size_t) ms * a.M + row) * ((size_t) a.njb * W) + col] = acc[s];
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

// Expected number of tail elements (beyond the SASO_NU a record holds) of one slot whose `halves` rows each get a
// Poisson(lambda) number of entries per chunk: sum over k >= SASO_NU of P(max > k).
static double saso_mean_tail(double lambda, int halves) {
    double pmf = exp(-lambda), cdf = pmf, mean = 0.0;
    for (int k = 1; k < 200; ++k) {
        if (k > SASO_NU) { double all = 1.0; for (int h = 0; h < halves; ++h) all *= cdf; mean += 1.0 - all; }   // cdf = F(k-1)
        pmf *= lambda / k; cdf += pmf;
    }
    return mean;
}

template <typename T>
int launch_saso_fast(const Ctr128& ctr, const Key64& key, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, int64_t ro, int64_t co,
                     int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                     T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    (void) S_cols;
    constexpr int W = SASO_ROWB / (int) sizeof(T);
    constexpr int HALVES = 32 / W;
    constexpr int NS = SASO_RW / HALVES;
    constexpr size_t SMEM_MAX = 227 * 1024 - 512;     // dynamic shared memory: the opt-in limit of a CTA less the kernels' static part (the tile kernel's mbarriers)
    // envelope of this kernel; everything else goes to the general path
    const bool cols_contig = (sx_l == 1 && (sx_j != 1 || N == 1));      // a single column is column-contiguous whatever its leading dimension says
    if (!cols_contig && sx_j != 1) return RBX_ERR_UNSUPPORTED;
    if (vec_nnz > 32 || M > SASO_MAX_RB * SASO_R || M < 1 || N < 1 || K < 64) return RBX_ERR_UNSUPPORTED;
    const int64_t ld = cols_contig ? sx_j : sx_l;
    // row-contiguous data is staged with 16-byte cp.async when its base and leading dimension are 16-byte aligned, element by
    // element otherwise; column-contiguous data always goes element by element (8 / 4-byte cp.async): nothing beyond the
    // natural alignment of T is required
    const bool x_elem = !cols_contig && (reinterpret_cast<uintptr_t>(X) % 16 != 0 || (ld * (int64_t) sizeof(T)) % 16 != 0);
    if (reinterpret_cast<uintptr_t>(X) % sizeof(T) != 0) return RBX_ERR_UNSUPPORTED;
    if (K > (int64_t) 0x7fffffff - 512 || N > 0x7fffffff) return RBX_ERR_UNSUPPORTED;
    if (alpha == (T) 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);

    // A CTA's 16 consumer warps own 64 rows each. A window of few rows would leave most of them idle and pile every chunk's
    // entries onto few accumulator slots (three per record, the rest in serial tails): d = 64 ran 14x slower than d = 1024 for
    // the same number of updates. Such windows are therefore replicated: R copies of the M rows, column cc of a chunk feeding
    // copy (cc mod R), so that the tile kernel sees Mv = R * M <= 1024 rows with 1 / R of the entries each; the copies are
    // summed by the reduce kernel together with the contraction splits (a fixed order: still bitwise reproducible).
    const int R = (M <= SASO_R / 2 && !getenv("RBX_SASO_NO_REPL")) ? (int)(SASO_R / M) : 1;
    const int64_t Mv = M * R;
    const int nbuckets = (int)((Mv + SASO_RW - 1) / SASO_RW);
    const int nrb = (nbuckets + SASO_WARPS - 1) / SASO_WARPS;           // row blocks of 1024 rows
    const int njb = (int)((N + W - 1) / W);
    // rows of A per chunk: as many as the ring stages (A chunk + the records and tails that go with it) fit in shared memory
    int CK = 0, cap = 0;
    const size_t nst = cols_contig ? SASO_NST_COLS : SASO_NST_ROWS;
    size_t stage = 0;
    for (int ck = cols_contig ? SASO_CK_MAX_COLS : SASO_CK_MAX; ck >= 64; ck -= (cols_contig ? 16 : 64)) {
        if ((size_t) ck * vec_nnz * 10 + (size_t)(2 * nbuckets * SASO_RW + 3 * nbuckets * NS + 2) * 4 + 16 > SMEM_MAX) continue;   // the prep kernel's sort
        // tail elements of one row block per chunk: mean + 8 sigma (Poisson tails; sigma^2 < 2 * mean); a bucket that
        // does not fit reads its tails from global memory
        const double lambda = (double) ck * (double) vec_nnz / ((double) S_rows * R);
        const double mean_rb = saso_mean_tail(lambda, HALVES) * SASO_WARPS * NS;
        int c = ((int)(mean_rb + 8.0 * sqrt(2.0 * mean_rb + 1.0)) + 32 + 31) & ~31;
        if (const char* ev = getenv("RBX_SASO_CAP")) { const int v = atoi(ev); if (v >= 0) c = (v + 31) & ~31; }    // tests: force the spill path
        const size_t pitch = SASO_ROWB + (cols_contig ? sizeof(T) : 0);
        const size_t sb = (((size_t) ck * pitch + pitch + 15) & ~(size_t) 15) + (size_t) SASO_WARPS * NS * HALVES * 16 + (size_t) c * HALVES * 4;
        if (nst * sb <= SMEM_MAX && (int64_t) ck <= ((K + 63) & ~(int64_t) 63)) { CK = ck; cap = c; stage = sb; break; }
    }
    if (CK == 0 || cap >= (1 << 19)) return RBX_ERR_UNSUPPORTED;
    const int nchunks = (int)((K + CK - 1) / CK);
    const int E = CK * (int) vec_nnz, Mp = nbuckets * SASO_RW, nslots = nbuckets * NS;
    const size_t prep_smem = (size_t) E * 10 + (size_t)(2 * Mp + 3 * nslots + 2) * 4 + 16;
    if (prep_smem > SMEM_MAX) return RBX_ERR_UNSUPPORTED;
    // split the contraction axis so that the grid is several waves of CTAs (one CTA per SM)
    int msplit = (6 * kNumSMs + njb * nrb - 1) / (njb * nrb);
    if (msplit > nchunks / 8) msplit = nchunks / 8;
    if (msplit < 1) msplit = 1;
    if (const char* ev = getenv("RBX_SASO_MSPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= nchunks) msplit = v; }

    const int stream_stride = ((nrb * cap + E) * HALVES + 3) & ~3;               // tail codes per chunk: regions, then spill
    const int rec_stride = nrb * SASO_WARPS * NS * HALVES * 4;                  // record words per chunk
    const int ws_stride = (nbuckets + 1 + 3) & ~3;
    uint32_t* codes = nullptr; uint32_t* rec = nullptr; uint32_t* ws = nullptr; T* partial = nullptr; int* flag = nullptr;
    const size_t Np = (size_t) njb * W;
    const bool direct = (msplit == 1 && R == 1);
    const size_t part_bytes = direct ? 0 : (size_t) msplit * Mv * Np * sizeof(T);
    if (ws_alloc(&codes, (size_t) nchunks * stream_stride * 4, stream) != cudaSuccess ||
        ws_alloc(&rec, (size_t) nchunks * rec_stride * 4, stream) != cudaSuccess ||
        ws_alloc(&ws, (size_t) nchunks * ws_stride * 4, stream) != cudaSuccess ||
        ws_alloc(&flag, sizeof(int), stream) != cudaSuccess ||
        (part_bytes && ws_alloc(&partial, part_bytes, stream) != cudaSuccess)) {
        cudaGetLastError();
        if (codes) cudaFreeAsync(codes, stream);
        if (rec) cudaFreeAsync(rec, stream);
        if (ws) cudaFreeAsync(ws, stream);
        if (flag) cudaFreeAsync(flag, stream);
        return RBX_ERR_ALLOC;
    }
    int rc = RBX_OK;
    if (cudaMemsetAsync(flag, 0, sizeof(int), stream) != cudaSuccess) { cudaGetLastError(); rc = RBX_ERR_CUDA; }
    if (rc == RBX_OK) {
        auto kern = vec_nnz <= 8 ? saso_prep_kernel<8, HALVES> : saso_prep_kernel<32, HALVES>;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) prep_smem);
        kern<<<nchunks, SASO_THREADS, prep_smem, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, CK, nbuckets, nrb, cols_contig ? 1 : 0, cap, R,
                                                           codes, stream_stride, rec, rec_stride, ws, ws_stride, flag);
        count_launch();
        rc = check_launch("saso_prep_kernel");
    }
    if (rc == RBX_OK) {
        SasoArgs<T> a;
        a.X = X; a.sx_l = sx_l; a.sx_j = sx_j; a.K = K; a.N = N;
        a.stream = codes; a.rec = rec; a.ws = ws;
        a.stream_stride = stream_stride; a.rec_stride = rec_stride; a.ws_stride = ws_stride;
        a.CK = CK; a.cap = cap;
        a.nbuckets = nbuckets; a.njb = njb; a.nrb = nrb; a.msplit = msplit; a.nchunks = nchunks;
        a.dbg_skip = getenv("RBX_SASO_SKIP") ? atoi(getenv("RBX_SASO_SKIP")) : 0;
        a.M = Mv; a.direct = direct ? 1 : 0; a.x_elem = x_elem ? 1 : 0;
        a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j; a.partial = partial; a.spill_flag = flag;
        const unsigned grid = (unsigned)(njb * msplit * nrb);
        const size_t smem = nst * stage;
        auto tile = cols_contig ? saso_tile_kernel<T, true, false> : saso_tile_kernel<T, false, false>;
        auto tile_spill = cols_contig ? saso_tile_kernel<T, true, true> : saso_tile_kernel<T, false, true>;
        cudaFuncSetAttribute(tile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        cudaFuncSetAttribute(tile_spill, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        tile<<<grid, SASO_TILE_THREADS, smem, stream>>>(a);
        tile_spill<<<grid, SASO_TILE_THREADS, smem, stream>>>(a);      // returns at once unless the prep kernel raised the flag
        count_launch(); count_launch();
        rc = check_launch("saso_tile_kernel");
        if (rc != RBX_OK) std::fprintf(stderr, "[rbx] saso_tile_kernel launch: grid %u, threads %d, dynamic smem %zu, CK %d, cap %d, M %lld, N %lld, K %lld, vec_nnz %lld, R %d, msplit %d\n",
                                       grid, (int) SASO_TILE_THREADS, smem, CK, cap, (long long) M, (long long) N, (long long) K, (long long) vec_nnz, R, msplit);
        if (rc == RBX_OK && !direct) {
            int64_t blocks = (M * N + 255) / 256;
            if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
            // partial[(ms * Mv + g * M + row) * Np + col] = part (ms * R + g) of M rows: msplit * R parts, added in that order
            saso_reduce_kernel<T><<<(unsigned) blocks, 256, 0, stream>>>(partial, msplit * R, M, N, (int64_t) Np, alpha, beta, C, sc_i, sc_j);
            count_launch();
            rc = check_launch("saso_reduce_kernel");
        }
    }
    cudaFreeAsync(codes, stream); cudaFreeAsync(rec, stream); cudaFreeAsync(ws, stream); cudaFreeAsync(flag, stream);
    if (partial) cudaFreeAsync(partial, stream);
    return rc;
}

template int launch_saso_fast<float>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, float,
                                     const float*, int64_t, int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_saso_fast<double>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, double,
                                      const double*, int64_t, int64_t, double, double*, int64_t, int64_t, cudaStream_t);

}  // namespace rbx
