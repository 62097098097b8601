// K3: sparse operator generation (SASO / LASO index + sign draws) on the device.
//
// Replaces sparse::repeated_fisher_yates (sparse_skops.hh:53-113), which is a serial loop on the CPU.
// The reference runs, for each minor-axis vector i, vec_nnz steps of a Fisher-Yates shuffle on a
// work array of length dim_major that it then restores to the identity (:93-103). Vectors are
// therefore independent: vector i, step j consumes Philox counter seed + i*vec_nnz + j (:78) and only
// words 0 (index) and 1 (sign) of the output. Here one thread owns one vector and keeps the shuffle
// lazily: only displaced positions are remembered, in a per-thread list of at most vec_nnz pairs (vec_nnz <= 16; longer
// vectors take one CTA each and a shared-memory hash table, below).
// Bit-exact by construction (integer arithmetic only).
#include <cstdlib>
#include "rbx_kernels.cuh"
#include "rbx_fy.cuh"

namespace rbx {

constexpr int kLazyMax = 128;   // longest vec_nnz handled by the per-thread lazy list

template <typename T, typename I, int KMAX>
__global__ void __launch_bounds__(128) fisher_yates_lazy_kernel(Ctr128 ctr, Key64 key, int vec_nnz, int64_t dim_major, int64_t dim_minor,
                                                               I* __restrict__ idxs_major, I* __restrict__ idxs_minor, T* __restrict__ vals) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= dim_minor) return;
    fy_vector<KMAX>(ctr, key, i, vec_nnz, dim_major, [&](int j, int64_t idx, bool negative) {
        const int64_t o = i * vec_nnz + j;
        idxs_major[o] = (I) idx;
        if (vals) vals[o] = negative ? (T) -1.0 : (T) 1.0;
        if (idxs_minor) idxs_minor[o] = (I) i;
    });
}

// 16 < vec_nnz <= kHashMax: one CTA per vector. The Philox draws of a vector are independent (counter seed + i * vec_nnz + j),
// only the shuffle is sequential, and most of its steps do not interact: step j reads positions j and ell_j >= j of the work
// array and writes ell_j, so it depends only on EARLIER steps that wrote j or ell_j. The work array is kept lazily in shared
// memory: positions below vec_nnz (every "j" lives there) in a direct array, positions from vec_nnz on in an open-addressing
// hash table that only ever holds displaced positions. The CTA takes 128 consecutive steps at a time, one per thread, and
// resolves them in rounds: every unresolved step enters its target into a small group-local table that keeps, per position,
// the earliest unresolved step writing it; a step is ready when it is that earliest writer of its own target and nobody
// unresolved writes its position j. Ready steps commute and run together; the earliest unresolved step is always ready, so
// the rounds end -- after one round when vec_nnz << dim_major, after two or three when the draws are dense (a full
// permutation). Two local tables alternate so that a round costs two barriers. Integer arithmetic only: the same indices as
// the serial loop, bit for bit.
constexpr int kHashMax = 8192;            // direct array 4 * 8192 + table 8 * 16384 + 8 KB of local tables = 168 KB of shared memory
constexpr int kHashGroup = 128;
constexpr int kLocalCap = 512;            // entries of a group-local table (4x the group)
constexpr uint32_t kHashEmpty = 0xffffffffu;

__device__ __forceinline__ uint32_t fy_hash(uint32_t e, int cap_log2) { return (e * 2654435761u) >> (32 - cap_log2); }

template <typename T, typename I>
__global__ void __launch_bounds__(kHashGroup) fisher_yates_hash_kernel(Ctr128 ctr, Key64 key, int k, uint32_t dim_major, int64_t dim_minor, int cap_log2,
                                                                      I* __restrict__ idxs_major, I* __restrict__ idxs_minor, T* __restrict__ vals) {
    extern __shared__ uint32_t fy_smem[];
    const uint32_t cap = 1u << cap_log2, mask = cap - 1, kpad = (uint32_t)((k + 31) & ~31);
    uint32_t* low = fy_smem;                              // low[p] = work[p] for p < k
    uint32_t* tkey = low + kpad;                          // displaced positions >= k: key, value
    uint32_t* tval = tkey + cap;
    uint32_t* lkey = tval + cap;                          // two group-local tables: position -> earliest unresolved step writing it
    uint32_t* lmin = lkey + 2 * kLocalCap;
    const uint32_t tid = threadIdx.x;

    auto find_or_insert = [&](uint32_t e) -> uint32_t {  // slot of key e; a new key starts as the identity (work[e] = e)
        uint32_t h = fy_hash(e, cap_log2);
        while (true) {
            const uint32_t prev = atomicCAS(&tkey[h], kHashEmpty, e);
            if (prev == kHashEmpty) { tval[h] = e; return h; }
            if (prev == e) return h;
            h = (h + 1) & mask;
        }
    };
    auto local_enter = [&](uint32_t* lk, uint32_t* lm, uint32_t p, uint32_t step) {
        uint32_t h = fy_hash(p, 9);
        while (true) {
            const uint32_t prev = atomicCAS(&lk[h], kHashEmpty, p);
            if (prev == kHashEmpty || prev == p) { atomicMin(&lm[h], step); return; }
            h = (h + 1) & (kLocalCap - 1);
        }
    };
    auto local_find = [&](const uint32_t* lk, const uint32_t* lm, uint32_t p) -> uint32_t {
        uint32_t h = fy_hash(p, 9);
        while (true) {
            const uint32_t kk = lk[h];
            if (kk == p) return lm[h];
            if (kk == kHashEmpty) return kHashEmpty;
            h = (h + 1) & (kLocalCap - 1);
        }
    };

    for (uint32_t t = tid; t < 2 * kLocalCap; t += kHashGroup) { lkey[t] = kHashEmpty; lmin[t] = kHashEmpty; }
    uint32_t parity = 0;
    for (int64_t i = blockIdx.x; i < dim_minor; i += gridDim.x) {
        __syncthreads();                                  // the previous vector's last steps are done with the arrays
        for (int t = tid; t < k; t += kHashGroup) low[t] = (uint32_t) t;
        for (uint32_t t = tid; t < cap; t += kHashGroup) tkey[t] = kHashEmpty;
        __syncthreads();
        for (int j0 = 0; j0 < k; j0 += kHashGroup) {
            const uint32_t j = (uint32_t) j0 + tid;
            const bool active = j < (uint32_t) k;
            uint32_t e = j;
            bool negative = false;
            if (active) {
                const uint4 rv = philox4x32_10(ctr_add(ctr, (uint64_t)(i * k + j)), key);
                e = j + rv.x % (dim_major - j);           // sparse_skops.hh:80
                negative = (rv.y & 1u) != 0;              // :87
            }
            const bool moves = active && e != j;          // e == j: the step reads work[j] and changes nothing
            bool unresolved = active;
            while (true) {
                uint32_t* lk = lkey + (parity & 1) * kLocalCap;
                uint32_t* lm = lmin + (parity & 1) * kLocalCap;
                if (unresolved && moves) local_enter(lk, lm, e, tid);
                __syncthreads();
                bool ready = false;
                if (unresolved) {
                    ready = local_find(lk, lm, j) == kHashEmpty;              // whoever writes position j is an earlier step
                    if (ready && moves) ready = local_find(lk, lm, e) == tid;
                }
                const int left = __syncthreads_or(unresolved && !ready);
                if (ready) {
                    const uint32_t at_j = low[j];
                    uint32_t out = at_j;
                    if (moves) {
                        if (e < (uint32_t) k) { out = low[e]; low[e] = at_j; }
                        else { const uint32_t h = find_or_insert(e); out = tval[h]; tval[h] = at_j; }
                    }
                    const int64_t o = i * k + j;
                    idxs_major[o] = (I) out;
                    if (vals) vals[o] = negative ? (T) -1.0 : (T) 1.0;
                    if (idxs_minor) idxs_minor[o] = (I) i;
                    unresolved = false;
                }
                // this round's table is read by nobody after the vote; it is used again two rounds (four barriers) from now
                for (uint32_t t = tid; t < kLocalCap; t += kHashGroup) { lk[t] = kHashEmpty; lm[t] = kHashEmpty; }
                ++parity;
                if (!left) break;
            }
        }
    }
}

// Fallback for vec_nnz > kHashMax (or a major axis beyond 32 bits): one CTA walks vectors with an explicit work array in global scratch
// (one array per CTA), exactly as the reference does, including the undo pass.
template <typename T, typename I>
__global__ void __launch_bounds__(32) fisher_yates_workarray_kernel(Ctr128 ctr, Key64 key, int64_t vec_nnz, int64_t dim_major, int64_t dim_minor,
                                                                   I* __restrict__ idxs_major, I* __restrict__ idxs_minor, T* __restrict__ vals,
                                                                   int64_t* __restrict__ scratch) {
    int64_t* work = scratch + (int64_t) blockIdx.x * (dim_major + vec_nnz);
    int64_t* pivots = work + dim_major;
    for (int64_t t = threadIdx.x; t < dim_major; t += blockDim.x) work[t] = t;
    __syncwarp();
    if (threadIdx.x != 0) return;
    for (int64_t i = blockIdx.x; i < dim_minor; i += gridDim.x) {
        for (int64_t j = 0; j < vec_nnz; ++j) {
            const uint4 rv = philox4x32_10(ctr_add(ctr, (uint64_t)(i * vec_nnz + j)), key);
            const int64_t ell = j + (int64_t)((uint64_t) rv.x % (uint64_t)(dim_major - j));
            pivots[j] = ell;
            const int64_t swap = work[ell];
            work[ell] = work[j];
            work[j] = swap;
            const int64_t o = i * vec_nnz + j;
            idxs_major[o] = (I) swap;
            if (vals) vals[o] = (rv.y % 2 == 0) ? (T) 1.0 : (T) -1.0;
            if (idxs_minor) idxs_minor[o] = (I) i;
        }
        for (int64_t j = vec_nnz - 1; j >= 0; --j) {   // restore the identity (sparse_skops.hh:93-103)
            const int64_t ell = pivots[j];
            const int64_t swap = work[ell];
            work[ell] = work[j];
            work[j] = swap;
        }
    }
}

template <typename T, typename I>
int launch_fisher_yates(const Ctr128& ctr, const Key64& key, int64_t vec_nnz, int64_t dim_major, int64_t dim_minor,
                        I* idxs_major, I* idxs_minor, T* vals, cudaStream_t stream) {
    if (vec_nnz > dim_major) return RBX_ERR_ARG;     // randblas_error_if(vec_nnz > dim_major), sparse_skops.hh:63
    if (vec_nnz <= 0 || dim_minor <= 0) return RBX_OK;
    int64_t hash_min = 16;        // measured cross-over of the two kernels (tools/micro/fy_threshold.py)
    if (const char* ev = getenv("RBX_FY_HASH_MIN")) hash_min = atoll(ev);
    if (vec_nnz > hash_min && vec_nnz <= kHashMax && dim_major <= 0xfffffffell && !getenv("RBX_FY_NO_HASH")) {
        int cap_log2 = 6;
        while ((1ll << cap_log2) < 2 * vec_nnz) ++cap_log2;
        const size_t kpad = (size_t)((vec_nnz + 31) & ~31ll), cap = (size_t) 1 << cap_log2;
        const size_t smem = 4 * (kpad + 2 * cap + 4 * kLocalCap);
        auto kern = fisher_yates_hash_kernel<T, I>;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem) != cudaSuccess) { cudaGetLastError(); return RBX_ERR_CUDA; }
        int per_sm = (int)((size_t) 200 * 1024 / (smem + 1024));
        if (per_sm > 12) per_sm = 12;
        if (per_sm < 1) per_sm = 1;
        const int64_t g = dim_minor < (int64_t) kNumSMs * per_sm ? dim_minor : (int64_t) kNumSMs * per_sm;
        kern<<<(unsigned) g, kHashGroup, smem, stream>>>(ctr, key, (int) vec_nnz, (uint32_t) dim_major, dim_minor, cap_log2, idxs_major, idxs_minor, vals);
        count_launch();
        return check_launch("fisher_yates_hash_kernel");
    }
    if (vec_nnz <= kLazyMax) {
        const unsigned g = (unsigned)((dim_minor + 127) / 128);
        if (vec_nnz <= 8)       fisher_yates_lazy_kernel<T, I, 8><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        else if (vec_nnz <= 32) fisher_yates_lazy_kernel<T, I, 32><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        else                    fisher_yates_lazy_kernel<T, I, kLazyMax><<<g, 128, 0, stream>>>(ctr, key, (int) vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals);
        count_launch();
        return check_launch("fisher_yates_lazy_kernel");
    }
    int64_t ctas = dim_minor < kNumSMs * 4 ? dim_minor : kNumSMs * 4;
    // one work array of (dim_major + vec_nnz) int64 per CTA: the CTA count follows the memory that is free (at most 1 GiB)
    const size_t per_cta = sizeof(int64_t) * (size_t)(dim_major + vec_nnz);
    const int64_t fit = (int64_t)(ws_budget((size_t) 1 << 30) / per_cta);
    if (fit < 1) return RBX_ERR_ALLOC;
    if (ctas > fit) ctas = fit;
    int64_t* scratch = nullptr;
    if (ws_alloc(&scratch, sizeof(int64_t) * (size_t)(ctas * (dim_major + vec_nnz)), stream) != cudaSuccess) {
        cudaGetLastError();
        return RBX_ERR_ALLOC;
    }
    fisher_yates_workarray_kernel<T, I><<<(unsigned) ctas, 32, 0, stream>>>(ctr, key, vec_nnz, dim_major, dim_minor, idxs_major, idxs_minor, vals, scratch);
    count_launch();
    int rc = check_launch("fisher_yates_workarray_kernel");
    cudaFreeAsync(scratch, stream);
    return rc;
}

#define RBX_INST_FY(T, I) \
    template int launch_fisher_yates<T, I>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, I*, I*, T*, cudaStream_t);
RBX_INST_FY(float, int32_t) RBX_INST_FY(float, int64_t) RBX_INST_FY(double, int32_t) RBX_INST_FY(double, int64_t)
#undef RBX_INST_FY

}  // namespace rbx
