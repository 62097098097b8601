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
//     (row, column) and emits, per warp, a stream of 32-bit codes (ring byte offset of the A row | sign << 31) plus a
//     count per accumulator slot. The main loop is a statically unrolled walk over the slots with a dynamic trip count
//     per slot: one broadcast LDS (code), one 128-byte gather LDS, one FP add. No atomics, no shared-memory
//     read-modify-write, fixed summation order => run-to-run identical results;
//   * double precision: the two half-warps work on different rows; the prep kernel pairs their streams and pads the
//     shorter with a code that points at a zero row of the ring;
//   * the row blocks (d / 1024) of one column segment are adjacent CTAs and re-read the A chunk through L2, so HBM
//     sees each byte of A once.
// Rejected after measurement: TMA boxes for the A chunk (256-byte box rows are TMA row-rate bound, ~28 cycles/row),
// cluster multicast (73 cycles/row), register tiles selected with brx.idx (~45 issue slots per select), B tile in
// shared memory with read-modify-write (latency bound at 16 warps/SM, 24 B of shared-memory traffic per update).
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
constexpr int SASO_NST = 3;         // ring stages
constexpr int SASO_PAD = 4;         // zero codes after each row block's stream (read-ahead of the main loop)
constexpr uint32_t SASO_SIGN = 0x80000000u;

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
// Outputs, per chunk:
//   stream[chunk * stream_stride + ...]  elements of HALVES codes; element e of slot q is at ws-relative position pos(q)+e
//   cnt[chunk * cnt_stride + q]          elements in slot q (q = bucket * NS + s; rows bucket*64 + h*NS + s, h < HALVES)
//   ws[chunk * ws_stride + b]            first element of bucket b (b = 0 .. nbuckets), row blocks start 16-byte aligned
//                                        and are followed by >= SASO_PAD zero elements
template <int KMAX, int HALVES>
__global__ void __launch_bounds__(SASO_THREADS) saso_prep_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t ro, int64_t co,
                                                               int64_t M, int64_t K, int CK, int nbuckets, int nrb, int cols_contig,
                                                               uint32_t* __restrict__ stream, int stream_stride,
                                                               uint16_t* __restrict__ cnt, int cnt_stride,
                                                               uint32_t* __restrict__ ws, int ws_stride) {
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
    uint16_t* trow = reinterpret_cast<uint16_t*>(slotoff + nslots + 1);   // E
    const int64_t chunk = blockIdx.x;
    // a code is the ring offset of A row c in elements: c * (row pitch / sizeof(T)); the pitch is 128 bytes, plus one
    // element when the data is column-contiguous (see saso_tile_kernel). Row CK of the ring is all zeros.
    const uint32_t pitch_el = 32u / HALVES + (cols_contig ? 1u : 0u);
    const uint32_t zero_code = (uint32_t) CK * pitch_el;

    uint32_t* so = stream + chunk * stream_stride;
    for (int e = t; e < stream_stride; e += SASO_THREADS) so[e] = zero_code;
    for (int r = t; r < Mp; r += SASO_THREADS) rowcnt[r] = 0;
    for (int e = t; e < E; e += SASO_THREADS) trow[e] = 0xFFFF;
    __syncthreads();
    const int64_t l = chunk * CK + t;
    if (t < CK && l < K) {
        const uint32_t c = (uint32_t) t;
        const uint32_t base = c * pitch_el;
        fy_vector<KMAX>(ctr, key, co + l, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
            const int64_t r = idx - ro;
            if (r >= 0 && r < M) {
                tcode[t * vec_nnz + j] = base | (negative ? SASO_SIGN : 0u);
                trow[t * vec_nnz + j] = (uint16_t) r;
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
    }
    __syncthreads();
    block_excl_scan(slotcnt, slotoff, nslots, wtot);
    // shift of each row block: aligned start, SASO_PAD zero elements after the previous block
    uint32_t shift[4] = { 0, 0, 0, 0 };
    for (int rb = 1; rb < nrb; ++rb) {
        const uint32_t raw = slotoff[rb * SASO_WARPS * NS];
        shift[rb] = ((raw + shift[rb - 1] + SASO_PAD + 3u) & ~3u) - raw;
    }
    uint16_t* co_ = cnt + chunk * cnt_stride;
    for (int q = t; q < cnt_stride; q += SASO_THREADS) co_[q] = q < nslots ? (uint16_t) slotcnt[q] : (uint16_t) 0;
    uint32_t* wo = ws + chunk * ws_stride;
    for (int b = t; b <= nbuckets; b += SASO_THREADS) {
        const int rb = b < nbuckets ? b / SASO_WARPS : nrb - 1;
        wo[b] = slotoff[b * NS] + shift[rb];
    }
    for (int q = t; q < nslots; q += SASO_THREADS) {
        const int b = q / NS, s = q % NS;
        const uint32_t n = slotcnt[q], p = slotoff[q] + shift[b / SASO_WARPS];
#pragma unroll
        for (int h = 0; h < HALVES; ++h) {
            const int row = b * SASO_RW + h * NS + s;
            const uint32_t nr = rowcnt[row];
            const uint32_t* src = sorted + rowoff[row];
            for (uint32_t i = 0; i < nr; ++i) so[(p + i) * HALVES + h] = src[i];
        }
        (void) n;
    }
}

template <typename T>
struct SasoArgs {
    const T* X; int64_t sx_l, sx_j; int64_t K, N;      // data view: X[l * sx_l + j * sx_j], one of the strides is 1
    const uint32_t* stream; const uint16_t* cnt; const uint32_t* ws;
    int stream_stride, cnt_stride, ws_stride;
    int CK, cap;                                       // rows per chunk; stream elements a ring stage can hold
    int nbuckets, njb, nrb, msplit, nchunks, dbg_skip;
    int64_t M;
    T alpha, beta;
    T* C; int64_t sc_i, sc_j;
    T* partial;                                        // [msplit][M][njb * W] when msplit > 1
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

// acc += (i < n) ? (+/-) *(T*)addr : 0, sign taken from bit 31 of code; a predicated load, no branch
template <typename T> __device__ __forceinline__ void gather_add_if(T& acc, uint32_t addr, uint32_t code, int n, int i);
template <> __device__ __forceinline__ void gather_add_if<double>(double& acc, uint32_t addr, uint32_t code, int n, int i) {
    unsigned long long bits;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.gt.s32 p, %2, %3;\n\tmov.b64 %0, 0;\n\t@p ld.shared.b64 %0, [%1];\n\t}"
                 : "=l"(bits) : "r"(addr), "r"(n), "r"(i));
    const double v = __longlong_as_double((long long) bits);
    acc += __hiloint2double(__double2hiint(v) ^ (int)(code & SASO_SIGN), __double2loint(v));
}
template <> __device__ __forceinline__ void gather_add_if<float>(float& acc, uint32_t addr, uint32_t code, int n, int i) {
    uint32_t bits;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.gt.s32 p, %2, %3;\n\tmov.b32 %0, 0;\n\t@p ld.shared.b32 %0, [%1];\n\t}"
                 : "=r"(bits) : "r"(addr), "r"(n), "r"(i));
    acc += __uint_as_float(bits ^ (code & SASO_SIGN));
}
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
// elements first .. n-1 of a slot whose codes sit in shared memory at code_addr, code_addr + step_bytes, ...:
// acc += (+/-) A value. A warp-uniform loop (bra.uni: no reconvergence bookkeeping).
template <typename T> __device__ __forceinline__ void slot_tail(T& acc, uint32_t code_addr, int first, int n, uint32_t lanebase, int step_bytes);
template <> __device__ __forceinline__ void slot_tail<double>(double& acc, uint32_t code_addr, int first, int n, uint32_t lanebase, int step_bytes) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 i, ca, code, a, lo, hi;\n\t.reg .f64 v;\n\t"
                 "setp.le.s32 p, %3, %2;\n\t@p bra.uni TAIL_DONE;\n\t"
                 "mov.b32 i, %2;\n\tmov.b32 ca, %1;\n"
                 "TAIL_LOOP:\n\t"
                 "ld.shared.b32 code, [ca];\n\t"
                 "shl.b32 a, code, 3;\n\tadd.u32 a, a, %4;\n\t"
                 "ld.shared.v2.b32 {lo, hi}, [a];\n\t"
                 "lop3.b32 hi, hi, code, 0x80000000, 0x78;\n\t"
                 "mov.b64 v, {lo, hi};\n\tadd.f64 %0, %0, v;\n\t"
                 "add.u32 ca, ca, %5;\n\tadd.s32 i, i, 1;\n\tsetp.lt.s32 p, i, %3;\n\t@p bra.uni TAIL_LOOP;\n"
                 "TAIL_DONE:\n\t}"
                 : "+d"(acc) : "r"(code_addr), "r"(first), "r"(n), "r"(lanebase), "r"(step_bytes));
}
template <> __device__ __forceinline__ void slot_tail<float>(float& acc, uint32_t code_addr, int first, int n, uint32_t lanebase, int step_bytes) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 i, ca, code, a, x;\n\t.reg .f32 v;\n\t"
                 "setp.le.s32 p, %3, %2;\n\t@p bra.uni TAIL_DONE;\n\t"
                 "mov.b32 i, %2;\n\tmov.b32 ca, %1;\n"
                 "TAIL_LOOP:\n\t"
                 "ld.shared.b32 code, [ca];\n\t"
                 "shl.b32 a, code, 2;\n\tadd.u32 a, a, %4;\n\t"
                 "ld.shared.b32 x, [a];\n\t"
                 "lop3.b32 x, x, code, 0x80000000, 0x78;\n\t"
                 "mov.b32 v, x;\n\tadd.f32 %0, %0, v;\n\t"
                 "add.u32 ca, ca, %5;\n\tadd.s32 i, i, 1;\n\tsetp.lt.s32 p, i, %3;\n\t@p bra.uni TAIL_LOOP;\n"
                 "TAIL_DONE:\n\t}"
                 : "+f"(acc) : "r"(code_addr), "r"(first), "r"(n), "r"(lanebase), "r"(step_bytes));
}
// volatile: keeps the code loads where they are written (one slot ahead of their use) instead of all up front
__device__ __forceinline__ uint32_t ld_code(const uint32_t* p, bool in_smem) {
    uint32_t v;
    if (in_smem) asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
    else asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

#ifndef SASO_NU_VAL
#define SASO_NU_VAL 3
#endif
constexpr int SASO_NU = SASO_NU_VAL;          // elements of a slot handled by predicated straight-line code; the rest loop

// slots [0, NS) of this warp: acc[s] += sum of the gathered (signed) A values of its elements.
// codes: this lane's first code (step = codes per element); lanebase: shared address of this lane's column in row 0 of
// the stage's A chunk. Address of an element = lanebase + (code << log2(sizeof(T))); the shift drops the sign bit.
template <typename T, int NS, bool IN_SMEM>
__device__ __forceinline__ void saso_consume(T (&acc)[NS], const uint32_t* codes, int step, uint32_t lanebase, const uint16_t* cnts) {
    constexpr int SH = sizeof(T) == 8 ? 3 : 2;
    uint32_t c[SASO_NU];
#pragma unroll
    for (int i = 0; i < SASO_NU; ++i) c[i] = ld_code(codes + i * step, IN_SMEM);
#pragma unroll
    for (int g = 0; g < NS; g += 8) {
        const uint4 cq = *reinterpret_cast<const uint4*>(cnts + g);
        const uint32_t cw[4] = { cq.x, cq.y, cq.z, cq.w };
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int n = (int)((cw[e >> 1] >> (16 * (e & 1))) & 0xffffu);
            uint32_t addr[SASO_NU], cc[SASO_NU];
#pragma unroll
            for (int i = 0; i < SASO_NU; ++i) { cc[i] = c[i]; addr[i] = lanebase + (c[i] << SH); }
            const uint32_t* cur = codes;
            codes += n * step;
#pragma unroll
            for (int i = 0; i < SASO_NU; ++i) c[i] = ld_code(codes + i * step, IN_SMEM);       // next slot's codes
#pragma unroll
            for (int i = 0; i < SASO_NU; ++i) gather_add_if<T>(acc[g + e], addr[i], cc[i], n, i);
            if (IN_SMEM) {
                slot_tail<T>(acc[g + e], smem_u32(cur + SASO_NU * step), SASO_NU, n, lanebase, step * 4);
            } else if (n > SASO_NU) {
#pragma unroll 1
                for (int i = SASO_NU; i < n; ++i) {
                    const uint32_t code = ld_code(cur + i * step, false);
                    acc[g + e] += lds_signed<T>(lanebase + (code << SH), code);
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
template <typename T, bool COLS_CONTIG>
__global__ void __launch_bounds__(SASO_THREADS, 1) saso_tile_kernel(const SasoArgs<T> a) {
    constexpr int W = SASO_ROWB / (int) sizeof(T);     // columns per CTA: 16 doubles / 32 floats
    constexpr int HALVES = 32 / W;                     // entries per warp instruction: 2 / 1
    constexpr int NS = SASO_RW / HALVES;               // accumulators per lane: 32 doubles / 64 floats
    constexpr int EP = 16 / (int) sizeof(T);           // elements per 16-byte piece
    constexpr int CNT_BYTES = SASO_WARPS * NS * 2;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t smem0 = smem_u32(smem);
    const int CK = a.CK;
    constexpr int PITCH = SASO_ROWB + (COLS_CONTIG ? (int) sizeof(T) : 0);
    const int a_bytes = CK * PITCH;
    const int az_bytes = (a_bytes + PITCH + 15) & ~15;                               // A chunk + its zero row
    const int stage_bytes = az_bytes + CNT_BYTES + a.cap * HALVES * 4;               // A | zero row | counts | codes

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
    const int b_first = rb * SASO_WARPS;
    const int b_last = (b_first + SASO_WARPS < a.nbuckets) ? b_first + SASO_WARPS : a.nbuckets;
    const int bucket = b_first + warp;                 // warp owns rows [bucket*64, bucket*64+64) of the window
    const bool active = bucket < a.nbuckets;

    if (tid < SASO_NST * (SASO_ROWB / 8))              // the zero rows of the stages
        *reinterpret_cast<uint2*>(smem + (size_t)(tid / (SASO_ROWB / 8)) * stage_bytes + a_bytes + (tid % (SASO_ROWB / 8)) * 8) = make_uint2(0, 0);

    // A chunk, slot counts and codes of chunk ch -> stage s; [lo, hi) = the row block's range of the code stream
    // tiles that touch neither edge of A take the copy loops without per-piece bounds arithmetic
    const bool cols_inside = j0 + W <= a.N;
    auto load_async = [&](int ch, int s, uint32_t lo, uint32_t hi) {
        const uint32_t st = smem0 + (uint32_t) s * stage_bytes;
        const bool inside = cols_inside && (int64_t)(ch + 1) * CK <= a.K;
        if (!COLS_CONTIG) {
            if (inside) {
                const T* src = a.X + ((int64_t) ch * CK + (tid >> 3)) * a.sx_l + j0 + (tid & 7) * EP;
                const int64_t step = (int64_t)(SASO_THREADS / 8) * a.sx_l;
                uint32_t dst = st + tid * 16;
                for (int i = 0; i < CK / (SASO_THREADS / 8); ++i) { cp_async16(dst, src, 16); dst += SASO_THREADS * 16; src += step; }
            } else {
                for (int p = tid; p < CK * 8; p += SASO_THREADS) {
                    const int row = p >> 3, q = p & 7;
                    const int64_t l = (int64_t) ch * CK + row, col = j0 + (int64_t) q * EP;
                    int64_t nb = (l < a.K) ? (a.N - col) * (int64_t) sizeof(T) : 0;
                    nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
                    const T* src = nb > 0 ? a.X + l * a.sx_l + col : a.X;
                    cp_async16(st + p * 16, src, (uint32_t) nb);
                }
            }
        } else if (tid < CK) {
            const int64_t l = (int64_t) ch * CK + tid;
            const uint32_t row = st + tid * PITCH;
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
        const uint8_t* csrc = reinterpret_cast<const uint8_t*>(a.cnt + (size_t) ch * a.cnt_stride + (size_t) rb * SASO_WARPS * NS);
        if (tid < CNT_BYTES / 16) cp_async16(st + az_bytes + tid * 16, csrc + tid * 16, 16);
        uint32_t nel = hi + SASO_PAD - lo;
        nel = nel > (uint32_t) a.cap ? (uint32_t) a.cap : nel;
        const int npiece = (int)((nel * HALVES * 4 + 15) / 16);
        const uint8_t* ssrc = reinterpret_cast<const uint8_t*>(a.stream + (size_t) ch * a.stream_stride + (size_t) lo * HALVES);
        const uint32_t sdst = st + az_bytes + CNT_BYTES;
        for (int p = tid; p < npiece; p += SASO_THREADS) cp_async16(sdst + p * 16, ssrc + p * 16, 16);
    };
    auto ws_at = [&](int ch, int b) -> uint32_t { return ch < ce ? __ldg(a.ws + (size_t) ch * a.ws_stride + b) : 0u; };

    constexpr int DIST = SASO_NST - 1;                 // prefetch distance in chunks
    for (int s = 0; s < DIST; ++s) {
        if (cb + s * cstep < ce) load_async(cb + s * cstep, s, ws_at(cb + s * cstep, b_first), ws_at(cb + s * cstep, b_last));
        cp_async_commit();
    }
    // stream positions, fetched one iteration ahead of their use
    uint32_t pf_lo = ws_at(cb + DIST * cstep, b_first), pf_hi = ws_at(cb + DIST * cstep, b_last);            // for the next prefetch
    uint32_t my_lo = ws_at(cb, b_first), my_first = 0, my_end = 0;                             // for the chunk being consumed
    if (active) { my_first = ws_at(cb, bucket); my_end = ws_at(cb, bucket + 1); }

    T acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) acc[s] = (T) 0;
    const uint32_t lx = (uint32_t) jl * (uint32_t) sizeof(T);

    for (int ch = cb, it = 0; ch < ce; ch += cstep, ++it) {
        cp_async_wait<DIST - 1>();
        __syncthreads();                                // chunk `it` has landed; everyone is done with chunk it-1
        const int nxt = ch + DIST * cstep;
        if (nxt < ce) load_async(nxt, (it + DIST) % SASO_NST, pf_lo, pf_hi);
        cp_async_commit();
        pf_lo = ws_at(nxt + cstep, b_first); pf_hi = ws_at(nxt + cstep, b_last);
        const uint32_t lo = my_lo, first = my_first, end = my_end;
        my_lo = ws_at(ch + cstep, b_first);
        if (active) { my_first = ws_at(ch + cstep, bucket); my_end = ws_at(ch + cstep, bucket + 1); }
        if (active && a.dbg_skip != 1) {
            const uint32_t st = smem0 + (uint32_t)(it % SASO_NST) * stage_bytes;
            const uint8_t* st8 = smem + (size_t)(it % SASO_NST) * stage_bytes;
            const uint16_t* cnts = reinterpret_cast<const uint16_t*>(st8 + az_bytes) + warp * NS;
            if (end + SASO_PAD - lo <= (uint32_t) a.cap) {
                const uint32_t* codes = reinterpret_cast<const uint32_t*>(st8 + az_bytes + CNT_BYTES) + (size_t)(first - lo) * HALVES + sub;
                saso_consume<T, NS, true>(acc, codes, HALVES, st + lx, cnts);
            } else {
                // the stage could not hold this warp's codes (far tail of the distribution): read them from global memory
                const uint32_t* codes = a.stream + (size_t) ch * a.stream_stride + (size_t) first * HALVES + sub;
                saso_consume<T, NS, false>(acc, codes, HALVES, st + lx, cnts);
            }
        }
    }
    cp_async_wait<0>();
    if (active) {
        const int64_t col = j0 + jl;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            const int64_t row = (int64_t) bucket * SASO_RW + sub * NS + s;
            if (row < a.M && col < a.N) {
                if (a.msplit == 1) {
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

template <typename T>
int launch_saso_fast(const Ctr128& ctr, const Key64& key, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, int64_t ro, int64_t co,
                     int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                     T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream) {
    (void) S_cols;
    constexpr int W = SASO_ROWB / (int) sizeof(T);
    constexpr int HALVES = 32 / W;
    constexpr int NS = SASO_RW / HALVES;
    constexpr size_t SMEM_MAX = 227 * 1024;
    // envelope of this kernel; everything else goes to the general path
    const bool cols_contig = (sx_l == 1 && sx_j != 1);
    if (!cols_contig && sx_j != 1) return RBX_ERR_UNSUPPORTED;
    if (vec_nnz > 32 || M > 4 * SASO_R || M < 1 || N < 1 || K < 64) return RBX_ERR_UNSUPPORTED;
    const int64_t ld = cols_contig ? sx_j : sx_l;
    if (reinterpret_cast<uintptr_t>(X) % 16 != 0 || (ld * (int64_t) sizeof(T)) % 16 != 0) return RBX_ERR_UNSUPPORTED;
    if (K > (int64_t) 0x7fffffff - 512 || N > 0x7fffffff) return RBX_ERR_UNSUPPORTED;
    if (alpha == (T) 0) return launch_scale_view<T>(C, M, N, sc_i, sc_j, beta, stream);

    const int nbuckets = (int)((M + SASO_RW - 1) / SASO_RW);
    const int nrb = (nbuckets + SASO_WARPS - 1) / SASO_WARPS;           // row blocks of 1024 rows
    const int njb = (int)((N + W - 1) / W);
    // rows of A per chunk: as many as three ring stages (A chunk + the codes that go with it) fit in shared memory
    int CK = 0, cap = 0;
    size_t stage = 0;
    for (int ck = 512; ck >= 64; ck -= 64) {
        const double ent_rb = (double) ck * (double) vec_nnz * (double)(M < SASO_R ? M : SASO_R) / (double) S_rows;
        const int c = (((int)(HALVES == 2 ? ent_rb : 1.25 * ent_rb) + 64) + 31) & ~31;
        const size_t pitch = SASO_ROWB + (cols_contig ? sizeof(T) : 0);
        const size_t sb = (((size_t) ck * pitch + pitch + 15) & ~(size_t) 15) + SASO_WARPS * NS * 2 + (size_t) c * HALVES * 4;
        if (SASO_NST * sb <= SMEM_MAX && (int64_t) ck <= ((K + 63) & ~(int64_t) 63)) { CK = ck; cap = c; stage = sb; break; }
    }
    if (CK == 0) return RBX_ERR_UNSUPPORTED;
    const int nchunks = (int)((K + CK - 1) / CK);
    const int E = CK * (int) vec_nnz, Mp = nbuckets * SASO_RW, nslots = nbuckets * NS;
    const size_t prep_smem = (size_t) E * 10 + (size_t)(2 * Mp + 2 * nslots + 2) * 4 + 16;
    if (prep_smem > SMEM_MAX) return RBX_ERR_UNSUPPORTED;
    // split the contraction axis so that the grid is several waves of CTAs (one CTA per SM)
    int msplit = (6 * kNumSMs + njb * nrb - 1) / (njb * nrb);
    if (msplit > nchunks / 8) msplit = nchunks / 8;
    if (msplit < 1) msplit = 1;
    if (const char* ev = getenv("RBX_SASO_MSPLIT")) { const int v = atoi(ev); if (v >= 1 && v <= nchunks) msplit = v; }

    const int stream_stride = ((E + nrb * 8 + SASO_PAD) * HALVES + 3) & ~3;      // codes per chunk
    const int cnt_stride = nrb * SASO_WARPS * NS;
    const int ws_stride = (nbuckets + 1 + 3) & ~3;
    uint32_t* codes = nullptr; uint16_t* cnt = nullptr; uint32_t* ws = nullptr; T* partial = nullptr;
    const size_t Np = (size_t) njb * W;
    const size_t part_bytes = msplit > 1 ? (size_t) msplit * M * Np * sizeof(T) : 0;
    if (cudaMallocAsync(&codes, (size_t) nchunks * stream_stride * 4, stream) != cudaSuccess ||
        cudaMallocAsync(&cnt, (size_t) nchunks * cnt_stride * 2, stream) != cudaSuccess ||
        cudaMallocAsync(&ws, (size_t) nchunks * ws_stride * 4, stream) != cudaSuccess ||
        (part_bytes && cudaMallocAsync(&partial, part_bytes, stream) != cudaSuccess)) {
        cudaGetLastError();
        if (codes) cudaFreeAsync(codes, stream);
        if (cnt) cudaFreeAsync(cnt, stream);
        if (ws) cudaFreeAsync(ws, stream);
        return RBX_ERR_ALLOC;
    }
    int rc = RBX_OK;
    {
        auto kern = vec_nnz <= 8 ? saso_prep_kernel<8, HALVES> : saso_prep_kernel<32, HALVES>;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) prep_smem);
        kern<<<nchunks, SASO_THREADS, prep_smem, stream>>>(ctr, key, (int) vec_nnz, S_rows, ro, co, M, K, CK, nbuckets, nrb, cols_contig ? 1 : 0,
                                                           codes, stream_stride, cnt, cnt_stride, ws, ws_stride);
        count_launch();
        rc = check_launch("saso_prep_kernel");
    }
    if (rc == RBX_OK) {
        SasoArgs<T> a;
        a.X = X; a.sx_l = sx_l; a.sx_j = sx_j; a.K = K; a.N = N;
        a.stream = codes; a.cnt = cnt; a.ws = ws;
        a.stream_stride = stream_stride; a.cnt_stride = cnt_stride; a.ws_stride = ws_stride;
        a.CK = CK; a.cap = cap;
        a.nbuckets = nbuckets; a.njb = njb; a.nrb = nrb; a.msplit = msplit; a.nchunks = nchunks;
        a.dbg_skip = getenv("RBX_SASO_SKIP") ? atoi(getenv("RBX_SASO_SKIP")) : 0;
        a.M = M; a.alpha = alpha; a.beta = beta; a.C = C; a.sc_i = sc_i; a.sc_j = sc_j; a.partial = partial;
        const unsigned grid = (unsigned)(njb * msplit * nrb);
        const size_t smem = SASO_NST * stage;
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
    cudaFreeAsync(codes, stream); cudaFreeAsync(cnt, stream); cudaFreeAsync(ws, stream);
    if (partial) cudaFreeAsync(partial, stream);
    return rc;
}

template int launch_saso_fast<float>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, float,
                                     const float*, int64_t, int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_saso_fast<double>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, double,
                                      const double*, int64_t, int64_t, double, double*, int64_t, int64_t, cudaStream_t);

}  // namespace rbx
