// Multi-GPU forms of the sketching path behind the C ABI: one process per GPU, an NCCL communicator owned
// by this library, and the row-sharded sketch -- every rank sketches its slab of the contraction axis with the operator's
// column offset advanced to the slab start (the operator is a pure function of (distribution, RNGState), so nothing about S
// is exchanged), then ONE collective sums the d x n partials over NVLink / NVSwitch: ncclReduceScatter when the caller wants
// its 1/world piece, ncclAllReduce in place otherwise. Column-sharded sketches need no code here: each rank calls the
// ordinary entry points on its own column slab.
//
// Reference contract: rtd/source/tutorial/updates.rst:59-80 ("sketching in pieces" through (ro_s, co_s)), skge.hh:182-189.
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): the library stays loadable on a box without NCCL, and inside a torch
// process the already-loaded NCCL (same soname) is the one used, so a process never holds two NCCL builds.
// The collective is not fused into the GEMM: at C5 the 268 MB reduce-scatter is ~1 ms of NVLink time behind a >= 18 ms
// per-rank sketch (8 GPUs, 65536-row slab) and it can only start when the last k-panel has been accumulated, so there is
// nothing to overlap tile by tile; it is issued on the sketch's stream right behind the last kernel.
#include <dlfcn.h>
#include <mutex>
#include <nccl.h>
#include "rbx_kernels.cuh"

extern "C" {
int rbx_sketch_dense_f32(char, char, char, char, int64_t, int64_t, int64_t, float, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                         const float*, char, int64_t, int64_t, const float*, int64_t, float, float*, int64_t, void*);
int rbx_sketch_dense_f64(char, char, char, char, int64_t, int64_t, int64_t, double, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                         const double*, char, int64_t, int64_t, const double*, int64_t, double, double*, int64_t, void*);
int rbx_sketch_sparseop_f32(char, char, char, char, int64_t, int64_t, int64_t, float, int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*,
                            int64_t, int64_t, const float*, int64_t, float, float*, int64_t, void*);
int rbx_sketch_sparseop_f64(char, char, char, char, int64_t, int64_t, int64_t, double, int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*,
                            int64_t, int64_t, const double*, int64_t, double, double*, int64_t, void*);
}

namespace rbx {

struct NcclApi {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclReduceScatter) ReduceScatter = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    bool ok = false;
};

static NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = { getenv("RBX_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so" };
        for (const char* nm : names) {
            if (!nm) continue;
            api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) { std::fprintf(stderr, "[rbx] NCCL not found (libnccl.so.2): %s\n", dlerror()); return; }
#define RBX_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.handle, "nccl" #name))
        RBX_SYM(GetUniqueId); RBX_SYM(CommInitRank); RBX_SYM(CommDestroy); RBX_SYM(ReduceScatter); RBX_SYM(AllReduce); RBX_SYM(GetErrorString);
        RBX_SYM(GetVersion);
#undef RBX_SYM
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.ReduceScatter && api.AllReduce && api.GetErrorString;
    });
    return api;
}

struct Comm {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
};

static int nccl_check(ncclResult_t r, const char* what) {
    if (r == ncclSuccess) return RBX_OK;
    std::fprintf(stderr, "[rbx] %s: %s\n", what, nccl().GetErrorString(r));
    return RBX_ERR_COMM;
}

template <typename T> struct NcclType;
template <> struct NcclType<float> { static constexpr ncclDataType_t v = ncclFloat; };
template <> struct NcclType<double> { static constexpr ncclDataType_t v = ncclDouble; };

template <typename T>
static int comm_sum(Comm* c, T* B, int64_t count, T* piece, cudaStream_t st) {
    if (!c || c->world == 1) {
        if (piece && piece != B && cudaMemcpyAsync(piece, B, sizeof(T) * (size_t) count, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return RBX_ERR_CUDA;
        return RBX_OK;
    }
    if (piece) {
        if (count % c->world != 0) return RBX_ERR_ARG;
        return nccl_check(nccl().ReduceScatter(B, piece, (size_t)(count / c->world), NcclType<T>::v, ncclSum, c->comm, st), "ncclReduceScatter");
    }
    return nccl_check(nccl().AllReduce(B, B, (size_t) count, NcclType<T>::v, ncclSum, c->comm, st), "ncclAllReduce");
}

// rank's slab of the contraction axis; B must be dense (ldb = its leading extent) because the collective runs over d * n
// contiguous elements
template <typename T, typename F>
static int shard_left_sketch(Comm* c, char layout, int64_t d, int64_t n, int64_t m_total, int64_t ldb, T beta, T* B, T* piece, int64_t align,
                             cudaStream_t st, F local) {
    if (d < 0 || n < 0 || m_total < 0) return RBX_ERR_ARG;
    if (ldb != (layout == 'C' ? d : n) || (layout != 'C' && layout != 'R')) return RBX_ERR_ARG;
    const int world = c ? c->world : 1, rank = c ? c->rank : 0;
    int64_t start = 0, len = 0;
    rbx_shard_range(m_total, world, rank, align, &start, &len);
    // beta * B enters the sum once (rank 0); a rank with no rows contributes beta_r * B
    const T my_beta = rank == 0 ? beta : (T) 0;
    int rc = local(start, len, my_beta);
    if (rc != RBX_OK) return rc;
    return comm_sum<T>(c, B, d * n, piece, st);
}

}  // namespace rbx

using namespace rbx;

extern "C" {

int rbx_comm_unique_id(uint8_t id[128]) {
    if (!nccl().ok) return RBX_ERR_COMM;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    return nccl_check(nccl().GetUniqueId(reinterpret_cast<ncclUniqueId*>(id)), "ncclGetUniqueId");
}

int rbx_comm_init(void** comm, int world, int rank, const uint8_t id[128]) {
    if (!comm || world < 1 || rank < 0 || rank >= world) return RBX_ERR_ARG;
    *comm = nullptr;
    if (!nccl().ok) return RBX_ERR_COMM;
    Comm* c = new Comm;
    c->world = world; c->rank = rank;
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    const int rc = nccl_check(nccl().CommInitRank(&c->comm, world, uid, rank), "ncclCommInitRank");
    if (rc != RBX_OK) { delete c; return rc; }
    *comm = c;
    return RBX_OK;
}

int rbx_comm_destroy(void* comm) {
    Comm* c = static_cast<Comm*>(comm);
    if (!c) return RBX_OK;
    const int rc = c->comm ? nccl_check(nccl().CommDestroy(c->comm), "ncclCommDestroy") : RBX_OK;
    delete c;
    return rc;
}

int rbx_comm_info(void* comm, int* world, int* rank, int* nccl_version) {
    Comm* c = static_cast<Comm*>(comm);
    if (world) *world = c ? c->world : 1;
    if (rank) *rank = c ? c->rank : 0;
    if (nccl_version) { int v = 0; if (nccl().ok && nccl().GetVersion) nccl().GetVersion(&v); *nccl_version = v; }
    return RBX_OK;
}

#define RBX_DEF_COMM(SUF, T)                                                                                                         \
int rbx_comm_reduce_scatter_##SUF(void* comm, T* B_dev, int64_t count, T* piece_dev, void* stream) {                                 \
    if (count < 0 || !piece_dev) return RBX_ERR_ARG;                                                                                 \
    return comm_sum<T>(static_cast<Comm*>(comm), B_dev, count, piece_dev, (cudaStream_t) stream);                                    \
}                                                                                                                                    \
int rbx_comm_all_reduce_##SUF(void* comm, T* B_dev, int64_t count, void* stream) {                                                   \
    if (count < 0) return RBX_ERR_ARG;                                                                                               \
    return comm_sum<T>(static_cast<Comm*>(comm), B_dev, count, (T*) nullptr, (cudaStream_t) stream);                                 \
}                                                                                                                                    \
int rbx_shard_sketch_dense_##SUF(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, T alpha,                  \
                                 int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4],                \
                                 const uint32_t key[2], int64_t ro_s, int64_t co_s, const T* A_local_dev, int64_t lda, T beta,       \
                                 T* B_dev, int64_t ldb, T* B_piece_dev, int64_t align, void* stream) {                               \
    return shard_left_sketch<T>(static_cast<Comm*>(comm), layout, d, n, m_total, ldb, beta, B_dev, B_piece_dev, align,               \
                                (cudaStream_t) stream, [&](int64_t start, int64_t len, T my_beta) {                                  \
        return rbx_sketch_dense_##SUF('L', layout, 'N', opA, d, n, len, alpha, D_rows, D_cols, family, major_axis, ctr, key,         \
                                      (const T*) nullptr, 'C', ro_s, co_s + start, A_local_dev, lda, my_beta, B_dev, ldb, stream); });\
}                                                                                                                                    \
int rbx_shard_sketch_sparseop_##SUF(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, T alpha,               \
                                    int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4],         \
                                    const uint32_t key[2], int64_t ro_s, int64_t co_s, const T* A_local_dev, int64_t lda, T beta,    \
                                    T* B_dev, int64_t ldb, T* B_piece_dev, int64_t align, void* stream) {                            \
    return shard_left_sketch<T>(static_cast<Comm*>(comm), layout, d, n, m_total, ldb, beta, B_dev, B_piece_dev, align,               \
                                (cudaStream_t) stream, [&](int64_t start, int64_t len, T my_beta) {                                  \
        return rbx_sketch_sparseop_##SUF('L', layout, 'N', opA, d, n, len, alpha, S_rows, S_cols, vec_nnz, major_axis, ctr, key,     \
                                         ro_s, co_s + start, A_local_dev, lda, my_beta, B_dev, ldb, stream); });                     \
}
RBX_DEF_COMM(f32, float)
RBX_DEF_COMM(f64, double)

}  // extern "C"
