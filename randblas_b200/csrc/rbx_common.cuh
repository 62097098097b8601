// Shared host/device helpers for the rbx kernels (error codes, launch checks, small index algebra).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <cuda_runtime.h>
#include "../../include/rbx.h"
#include "rbx_rng.cuh"

namespace rbx {

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids are sized in multiples of this

inline int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        std::fprintf(stderr, "[rbx] %s: %s\n", what, cudaGetErrorString(e));
        return RBX_ERR_CUDA;
    }
    return RBX_OK;
}
// Work threshold (multiply-adds) below which the dense sketch takes the one-launch SIMT kernel instead of the tensor-core
// kernels. RBX_DENSE_MIN_MACS overrides it: the randomized parity tests set it to 0 so that small ragged shapes reach the
// tensor-core kernels too.
inline double dense_min_macs(double dflt) {
    const char* ev = getenv("RBX_DENSE_MIN_MACS");
    return ev ? atof(ev) : dflt;
}
// Stream-ordered workspaces come from a PRIVATE memory pool per device (cudaMemPoolCreate), not from the device's default
// pool that the host application and other libraries (torch's allocator, cuBLAS) share: nothing this library does changes
// their pool's behaviour. The private pool keeps up to a bounded amount mapped between calls (release threshold = 1/8 of
// the device's memory, 22 GiB on a B200, and only ever what a call really used; RBX_WS_KEEP_MB overrides) -- a workspace freed by one call is still mapped for the
// next, because re-mapping a few hundred MB costs more than the kernels here -- and hands anything above that back to the
// driver at the next synchronisation. rbx_trim_workspace() returns all of it. Creation is guarded by std::call_once.
struct WsState {
    cudaMemPool_t pool[64] = {};
    cudaStream_t aux[64] = {};
    std::once_flag pool_once[64], aux_once[64];
};
inline WsState& ws_state() { static WsState s; return s; }
inline cudaMemPool_t ws_pool(int dev) {
    WsState& S = ws_state();
    if (dev < 0 || dev >= 64) return nullptr;
    std::call_once(S.pool_once[dev], [&S, dev] {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) { cudaGetLastError(); return; }
        size_t free_b = 0, total_b = 0;
        unsigned long long keep = 4ull << 30;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) keep = (unsigned long long) total_b / 8; else cudaGetLastError();
        if (const char* ev = getenv("RBX_WS_KEEP_MB")) keep = (unsigned long long) atoll(ev) << 20;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        S.pool[dev] = pool;
    });
    return S.pool[dev];
}
inline cudaError_t ws_alloc(void** p, size_t bytes, cudaStream_t stream) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return cudaGetLastError();
    cudaMemPool_t pool = ws_pool(dev);
    return pool ? cudaMallocFromPoolAsync(p, bytes, pool, stream) : cudaMallocAsync(p, bytes, stream);
}
template <typename P> inline cudaError_t ws_alloc(P** p, size_t bytes, cudaStream_t stream) {
    return ws_alloc(reinterpret_cast<void**>(p), bytes, stream);
}
// bytes a single workspace may take: a quarter of what is free now (the sketch_sparse operator slab sizes itself with this)
inline size_t ws_budget(size_t want) {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return want; }
    int dev = 0; cudaGetDevice(&dev);
    unsigned long long cached = 0, used = 0;                   // memory the private pool holds but no call is using counts as free
    if (cudaMemPool_t pool = ws_pool(dev)) {
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &cached) != cudaSuccess) { cudaGetLastError(); cached = 0; }
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) != cudaSuccess) { cudaGetLastError(); used = 0; }
    }
    const size_t avail = (free_b + (size_t)(cached > used ? cached - used : 0)) / 4;
    return want < avail ? want : avail;
}

// one auxiliary (non-blocking) stream per device: operator panels are generated on it while the previous panel's GEMM runs
inline cudaStream_t aux_stream() {
    WsState& S = ws_state();
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::call_once(S.aux_once[dev], [&S, dev] {
        if (cudaStreamCreateWithFlags(&S.aux[dev], cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); S.aux[dev] = nullptr; }
    });
    return S.aux[dev];
}

#define RBX_CUDA_TRY(expr)                                                                   \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            std::fprintf(stderr, "[rbx] %s failed: %s\n", #expr, cudaGetErrorString(e__));   \
            return RBX_ERR_CUDA;                                                             \
        }                                                                                    \
    } while (0)

inline Ctr128 mk_ctr(const uint32_t* c) { Ctr128 r; for (int i = 0; i < 4; ++i) r.v[i] = c[i]; return r; }
inline Key64  mk_key(const uint32_t* k) { Key64 r; r.v[0] = k[0]; r.v[1] = k[1]; return r; }
inline void   put_ctr(const Ctr128& c, uint32_t* out) { if (out) for (int i = 0; i < 4; ++i) out[i] = c.v[i]; }

// dist_to_layout / major_axis_length (dense_skops.hh:297-316)
inline bool natural_is_rowmajor(int64_t n_rows, int64_t n_cols, char major_axis) {
    const bool wide = n_rows < n_cols, lng = (major_axis == 'L');
    return wide ? lng : !lng;
}
inline int64_t major_axis_length(int64_t n_rows, int64_t n_cols, char major_axis) {
    const int64_t mx = n_rows > n_cols ? n_rows : n_cols, mn = n_rows < n_cols ? n_rows : n_cols;
    return major_axis == 'L' ? mx : mn;
}

// A dense-distribution operator window as the kernels see it: entry (i, j) of op-free S[ro+i, co+j]
// is lane ((pc0 + pj) & 3) of generate4(philox(ctr + (pr0 + pi) * W + (pc0 + pj) / 4)), where the
// parent is S when the natural layout is RowMajor (pi = i, pj = j) and S^T otherwise (pi = j, pj = i).
// dense_skops.hh:109-131,491-503.
struct DenseOpWindow {
    Ctr128 ctr; Key64 key;
    int64_t W;          // counters per parent row = ceil(ma_len / 4)
    int64_t pr0, pc0;   // window origin in parent coordinates
    int parent_is_S;    // 1: parent rows are rows of S; 0: parent rows are columns of S
    float scale;        // 1 for Gaussian, (float)sqrt(3) for uniform (applied in T precision by callers)
    char family;
};

inline int make_dense_window(int64_t D_rows, int64_t D_cols, char family, char major_axis,
                             const uint32_t* ctr, const uint32_t* key, int64_t ro_s, int64_t co_s,
                             DenseOpWindow* w) {
    if (family != 'G' && family != 'U') return RBX_ERR_ARG;
    if (major_axis != 'L' && major_axis != 'S') return RBX_ERR_ARG;
    const int64_t ma = major_axis_length(D_rows, D_cols, major_axis);
    w->ctr = mk_ctr(ctr); w->key = mk_key(key);
    w->W = (ma + 3) / 4;
    w->parent_is_S = natural_is_rowmajor(D_rows, D_cols, major_axis) ? 1 : 0;
    w->pr0 = w->parent_is_S ? ro_s : co_s;
    w->pc0 = w->parent_is_S ? co_s : ro_s;
    w->family = family;
    w->scale = 1.0f;
    return RBX_OK;
}

}  // namespace rbx
