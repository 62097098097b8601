// RandBLAS_b200.hh -- the RandBLAS C++ surface for the sketching hot path, implemented on the rbx C ABI (include/rbx.h).
//
// A caller of RandBLAS (https://github.com/BallisticLA/RandBLAS) that uses the types and functions below can include
// this header instead of <RandBLAS.hh> and link librbx.so: same namespaces, names, template parameters, argument order
// and precondition failures (RandBLAS::Error, like RandBLAS::exceptions::Error). Bodies marshal plain data to the
// extern "C" entry points; no CUDA header is needed to compile a caller (g++ is enough).
//
// Reference interfaces mirrored (paths under the RandBLAS tree):
//   RNGState                      base.hh:161-232          MajorAxis                 base.hh:138-150
//   DenseDist / DenseSkOp         dense_skops.hh:204-419   fill_dense x3             dense_skops.hh:486-592
//   SparseDist / SparseSkOp       sparse_skops.hh:134-377  fill_sparse               sparse_skops.hh:389-413
//   COOMatrix/CSRMatrix/CSCMatrix sparse_data/*.hh         sketch_general x4         skge.hh:771-1214
//   sketch_vector x2              skve.hh:152-258          sketch_symmetric x4       sksy.hh:165-537
//   sketch_sparse x2              sksp.hh:464-616
//
// Memory: by default pointers are HOST pointers, exactly like the reference; the call stages through HBM
// (rbx_host_* entry points). Define RBX_DEVICE_POINTERS before including this header (or set
// RandBLAS::device_pointers() = true at run time) when A, B and operator buffers already live in device memory.
// Only RNG = r123::Philox4x32 (the reference's default) is supported.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <limits>
#include <vector>
#include <algorithm>
#include <cmath>
#include <new>
#include <ostream>
#include "rbx.h"

#ifndef BLAS_HH      // BLAS++'s enums, when BLAS++ itself is not on the include path
namespace blas {
enum class Layout : char { ColMajor = 'C', RowMajor = 'R' };
enum class Op : char { NoTrans = 'N', Trans = 'T', ConjTrans = 'C' };
enum class Uplo : char { Upper = 'U', Lower = 'L', General = 'G' };
enum class Side : char { Left = 'L', Right = 'R' };
}
#endif

// The two Random123 array types RandBLAS exposes through RNGState (Random123 array.h: `v`, operator[], incr(n) with
// carry, STL-style access) and the generator tag. Only what RandBLAS callers touch is mirrored; the generator itself
// runs on the device (rbx_rng.cuh).
namespace r123 {
template <int N> struct ArrayNx32 {
    using value_type = uint32_t;
    enum { static_size = N };
    uint32_t v[N];
    uint32_t& operator[](size_t i) { return v[i]; }
    const uint32_t& operator[](size_t i) const { return v[i]; }
    static constexpr size_t size() { return N; }
    uint32_t* begin() { return v; }
    uint32_t* end() { return v + N; }
    const uint32_t* begin() const { return v; }
    const uint32_t* end() const { return v + N; }
    uint32_t* data() { return v; }
    const uint32_t* data() const { return v; }
    // add a 64-bit n to the little-endian N x 32-bit integer, with carry (pinned by test_r123.cc:679-766)
    ArrayNx32& incr(unsigned long long n = 1) {
        uint64_t carry = n;
        for (int i = 0; i < N && carry; ++i) { const uint64_t t = (uint64_t) v[i] + (carry & 0xFFFFFFFFull); v[i] = (uint32_t) t; carry = (carry >> 32) + (t >> 32); }
        return *this;
    }
    bool operator==(const ArrayNx32& o) const { return std::memcmp(v, o.v, sizeof v) == 0; }
    bool operator!=(const ArrayNx32& o) const { return !(*this == o); }
};
using Array4x32 = ArrayNx32<4>;
using Array2x32 = ArrayNx32<2>;
struct Philox4x32 { using ctr_type = Array4x32; using key_type = Array2x32; };
}  // namespace r123

namespace RandBLAS {

class Error : public std::runtime_error {
public:
    explicit Error(const std::string& what) : std::runtime_error(what) {}
};
namespace exceptions { using Error = ::RandBLAS::Error; }

#define randblas_require(cond) do { if (!(cond)) throw ::RandBLAS::Error(std::string("(" #cond ") was required, but did not hold, in function ") + __func__); } while (0)

inline bool& device_pointers() {
#ifdef RBX_DEVICE_POINTERS
    static bool v = true;
#else
    static bool v = false;
#endif
    return v;
}
// stream used by the device-pointer entry points of the calling thread (a cudaStream_t; nullptr = default stream)
inline void*& current_stream() { static thread_local void* s = nullptr; return s; }

namespace detail {
inline void check(int rc, const char* where) {
    if (rc == RBX_OK) return;
    if (rc == RBX_ERR_ARG) throw Error(std::string("a precondition of ") + where + " was required, but did not hold");
    throw std::runtime_error(std::string(where) + ": " + rbx_strerror(rc));
}
inline char ch(blas::Layout l) { return l == blas::Layout::ColMajor ? 'C' : 'R'; }
inline char ch(blas::Op o) { return o == blas::Op::NoTrans ? 'N' : 'T'; }
template <typename T> struct is_real : std::integral_constant<bool, std::is_same<T, float>::value || std::is_same<T, double>::value> {};

// RAII device buffer for the host-pointer paths that have no single C entry point
struct DevBuf {
    void* p = nullptr;
    DevBuf() = default;
    DevBuf(const void* host, uint64_t bytes) { upload(host, bytes); }
    void alloc(uint64_t bytes) { check(rbx_malloc(&p, bytes), "rbx_malloc"); }
    void upload(const void* host, uint64_t bytes) { alloc(bytes); if (host) check(rbx_memcpy_h2d(p, host, bytes, nullptr), "rbx_memcpy_h2d"); }
    ~DevBuf() { rbx_free(p); }
    DevBuf(const DevBuf&) = delete; DevBuf& operator=(const DevBuf&) = delete;
};
}  // namespace detail

// ---------------------------------------------------------------------------------------------------
// base.hh
// ---------------------------------------------------------------------------------------------------
enum class MajorAxis : char { Short = 'S', Long = 'L', Undefined = 'U' };

template <typename RNG = r123::Philox4x32>
struct RNGState {
    static_assert(std::is_same<RNG, r123::Philox4x32>::value, "only r123::Philox4x32 is implemented on the device");
    using generator = RNG;
    using ctr_type = typename RNG::ctr_type;
    using key_type = typename RNG::key_type;
    using ctr_uint = typename ctr_type::value_type;
    using key_uint = typename key_type::value_type;
    const static int len_c = ctr_type::static_size;
    const static int len_k = key_type::static_size;
    ctr_type counter;                          // state.counter.v, state.counter.incr(n), state.counter.v[i] as in the reference
    key_type key;
    RNGState() : counter{{0, 0, 0, 0}}, key{{0, 0}} {}
    RNGState(key_type const& k) : counter{{0, 0, 0, 0}}, key(k) {}
    RNGState(ctr_type const& c, key_type const& k) : counter(c), key(k) {}
    RNGState(key_uint k) : counter{{0, 0, 0, 0}}, key{{k, 0}} {}       // base.hh:205
    // base.hh:217-232: plain copies
};
template <typename RNG>
inline std::ostream& operator<<(std::ostream& out, const RNGState<RNG>& s) {           // base.hh:234-251
    out << "counter : {";
    for (int i = 0; i < s.len_c - 1; ++i) out << s.counter[i] << ", ";
    out << s.counter[s.len_c - 1] << "}\n" << "key     : {";
    for (int i = 0; i < s.len_k - 1; ++i) out << s.key[i] << ", ";
    out << s.key[s.len_k - 1] << "}";
    return out;
}

inline int64_t safe_int_product(int64_t a, int64_t b) {                 // base.hh:121-135
    if (a != 0 && b != 0) {
        const int64_t c = (int64_t)((uint64_t) a * (uint64_t) b);
        if (c / a != b) throw std::overflow_error("overflow in integer product");
        return c;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// dense_skops.hh
// ---------------------------------------------------------------------------------------------------
enum class DenseDistName : char { Gaussian = 'G', Uniform = 'U', BlackBox = 'B' };

struct DenseDist {
    const int64_t n_rows;
    const int64_t n_cols;
    const DenseDistName family;
    const MajorAxis major_axis;
    DenseDist(int64_t n_rows, int64_t n_cols, DenseDistName dn = DenseDistName::Gaussian, MajorAxis ma = MajorAxis::Long)
        : n_rows(n_rows), n_cols(n_cols), family(dn), major_axis(ma) {}
};

inline blas::Layout dist_to_layout(const DenseDist& D) {                 // dense_skops.hh:297-310
    randblas_require(D.major_axis != MajorAxis::Undefined);
    return rbx_dist_to_layout(D.n_rows, D.n_cols, (char) D.major_axis) == 'C' ? blas::Layout::ColMajor : blas::Layout::RowMajor;
}
inline int64_t major_axis_length(const DenseDist& D) {                   // dense_skops.hh:312-316
    randblas_require(D.major_axis != MajorAxis::Undefined);
    return rbx_major_axis_length(D.n_rows, D.n_cols, (char) D.major_axis);
}
template <typename T> inline T isometry_scale_factor(DenseDist D) { return std::pow((T) std::min(D.n_rows, D.n_cols), (T) -0.5); }

namespace dense {
template <typename RNG>
inline RNGState<RNG> compute_next_state(DenseDist D, RNGState<RNG> state) {   // dense_skops.hh:172-191
    if (D.major_axis == MajorAxis::Undefined) return state;
    RNGState<RNG> out = state;
    rbx_dense_next_state(D.n_rows, D.n_cols, (char) D.major_axis, state.counter.v, out.counter.v);
    return out;
}
}  // namespace dense

template <typename T, typename RNG = r123::Philox4x32>
struct DenseSkOp {
    using generator = RNG;
    using state_type = RNGState<RNG>;
    using buffer_type = T;
    const DenseDist dist;
    const RNGState<RNG> seed_state;
    const RNGState<RNG> next_state;
    T* buff = nullptr;                         // host memory (or device memory when device_pointers())
    blas::Layout layout;
    bool del_buff_on_destruct = false;
    const int64_t n_rows, n_cols;

    DenseSkOp(DenseDist dist, RNGState<RNG> const& state, T* buff = nullptr)
        : dist(dist), seed_state(state), next_state(dense::compute_next_state(dist, state)), buff(buff),
          layout(dist_to_layout(dist)), n_rows(dist.n_rows), n_cols(dist.n_cols) {}
    DenseSkOp(DenseDist dist, uint32_t key, T* buff = nullptr) : DenseSkOp(dist, RNGState<RNG>(key), buff) {}
    DenseSkOp(DenseDist dist, RNGState<RNG> const& seed, RNGState<RNG> const& next, T* buff, blas::Layout layout, bool del)
        : dist(dist), seed_state(seed), next_state(next), buff(buff), layout(layout), del_buff_on_destruct(del),
          n_rows(dist.n_rows), n_cols(dist.n_cols) {}
    ~DenseSkOp() {
        if (del_buff_on_destruct && buff) { if (device_pointers()) rbx_free(buff); else delete[] buff; }
    }
    DenseSkOp(const DenseSkOp&) = delete;
};

// ---------------------------------------------------------------------------------------------------
// sparse_skops.hh
// ---------------------------------------------------------------------------------------------------
struct SparseDist {
    const int64_t n_rows;
    const int64_t n_cols;
    const int64_t vec_nnz;
    const MajorAxis major_axis = MajorAxis::Short;
};

namespace sparse {
template <typename RNG>
inline RNGState<RNG> compute_next_state(SparseDist D, RNGState<RNG> state) {  // sparse_skops.hh:115-126
    RNGState<RNG> out = state;
    rbx_sparse_next_state(D.n_rows, D.n_cols, D.vec_nnz, (char) D.major_axis, state.counter.v, out.counter.v);
    return out;
}
}  // namespace sparse

template <typename T, typename RNG = r123::Philox4x32, typename sint_t = int64_t>
struct SparseSkOp {
    static_assert(std::is_same<sint_t, int64_t>::value || std::is_same<sint_t, int32_t>::value, "sint_t must be int64_t or int");
    using generator = RNG;
    using state_type = RNGState<RNG>;
    using index_t = sint_t;
    using scalar_t = T;
    const SparseDist dist;
    const RNGState<RNG> seed_state;
    const RNGState<RNG> next_state;
    const bool own_memory = true;
    bool known_filled = false;
    sint_t* rows = nullptr;
    sint_t* cols = nullptr;
    T* vals = nullptr;
    const int64_t n_rows, n_cols;

    SparseSkOp(SparseDist dist, const RNGState<RNG>& state)              // sparse_skops.hh:261-279
        : dist(dist), seed_state(state), next_state(sparse::compute_next_state(dist, state)), n_rows(dist.n_rows), n_cols(dist.n_cols) {
        const int64_t minor = (dist.major_axis == MajorAxis::Short) ? std::max(dist.n_rows, dist.n_cols) : std::min(dist.n_rows, dist.n_cols);
        const int64_t nnz = dist.vec_nnz * minor;
        // the owning constructor allocates where the caller's pointers live: host memory like the reference, or device memory
        // when device_pointers() is set (fill_sparse then writes the triplets in place)
        on_device = device_pointers();
        auto get = [&](size_t bytes) -> void* {
            if (!on_device) return ::operator new[](bytes);
            void* p = nullptr; detail::check(rbx_malloc(&p, bytes), "rbx_malloc"); return p;
        };
        rows = (sint_t*) get(nnz * sizeof(sint_t)); cols = (sint_t*) get(nnz * sizeof(sint_t)); vals = (T*) get(nnz * sizeof(T));
    }
    SparseSkOp(SparseDist dist, uint32_t key) : SparseSkOp(dist, RNGState<RNG>(key)) {}
    // views of caller-owned triplets (sparse_skops.hh:266-290); known_filled = true says they already hold the operator
    SparseSkOp(SparseDist dist, const RNGState<RNG>& state, sint_t* rows, sint_t* cols, T* vals, bool known_filled = true)
        : dist(dist), seed_state(state), next_state(sparse::compute_next_state(dist, state)), own_memory(false),
          known_filled(known_filled), rows(rows), cols(cols), vals(vals), n_rows(dist.n_rows), n_cols(dist.n_cols) {
        randblas_require(dist.n_rows > 0); randblas_require(dist.n_cols > 0); randblas_require(dist.vec_nnz > 0);
    }
    // shallow copies, possibly transposed (sparse_skops.hh:293-310; the reference's self-assignment leaves known_filled
    // false there, which transpose() below corrects on the copy it returns)
    SparseSkOp(SparseDist dist, const RNGState<RNG>& seed_state, sint_t* rows, sint_t* cols, T* vals, const RNGState<RNG>& next_state)
        : dist(dist), seed_state(seed_state), next_state(next_state), own_memory(false), rows(rows), cols(cols), vals(vals),
          n_rows(dist.n_rows), n_cols(dist.n_cols) {
        randblas_require(dist.n_rows > 0); randblas_require(dist.n_cols > 0); randblas_require(dist.vec_nnz > 0);
    }
    SparseSkOp(SparseDist dist, uint32_t key, sint_t* rows, sint_t* cols, T* vals) : SparseSkOp(dist, RNGState<RNG>(key), rows, cols, vals) {}
    ~SparseSkOp() {
        if (!own_memory) return;
        for (void* p : { (void*) rows, (void*) cols, (void*) vals }) { if (!p) continue; if (on_device) rbx_free(p); else ::operator delete[](p); }
    }
    bool on_device = false;
    SparseSkOp(const SparseSkOp&) = delete;
    SparseSkOp(SparseSkOp&& o)                                            // sparse_skops.hh:362-376
        : dist(o.dist), seed_state(o.seed_state), next_state(o.next_state), own_memory(o.own_memory), known_filled(o.known_filled),
          rows(o.rows), cols(o.cols), vals(o.vals), n_rows(o.n_rows), n_cols(o.n_cols), on_device(o.on_device) { o.rows = nullptr; o.cols = nullptr; o.vals = nullptr; }
};

template <typename SKOP>
inline int64_t nnz(const SKOP& S) {                                       // sparse_skops.hh:465-472
    const bool saso = S.dist.major_axis == MajorAxis::Short;
    const int64_t minor = saso ? std::max(S.dist.n_rows, S.dist.n_cols) : std::min(S.dist.n_rows, S.dist.n_cols);
    return S.dist.vec_nnz * minor;
}
template <typename SparseSkOpT> inline void fill_sparse(SparseSkOpT& S);

// ---------------------------------------------------------------------------------------------------
// sparse_data/{coo,csr,csc}_matrix.hh -- own-or-view containers (index_base Zero only)
// ---------------------------------------------------------------------------------------------------
namespace sparse_data {
enum class IndexBase : char { Zero = 'Z', One = 'O' };

enum class NonzeroSort : char { CSC = 'c', CSR = 'r', None = 'n' };

namespace detail_sp {
// arrays of an owning container live where the caller's pointers live: host memory by default (new[] / delete[], like the
// reference), device memory when device_pointers() was set at reserve() time
template <typename U> inline U* alloc_(int64_t n, bool on_device) {
    if (!on_device) return new U[n];
    void* p = nullptr;
    if (rbx_malloc(&p, (uint64_t) n * sizeof(U)) != RBX_OK) throw std::bad_alloc();
    return static_cast<U*>(p);
}
template <typename U> inline void free_(U* p, bool on_device) { if (!p) return; if (on_device) rbx_free(p); else delete[] p; }
}  // namespace detail_sp

template <typename T, typename sint_t = int64_t>
struct COOMatrix {
    using scalar_t = T; using index_t = sint_t;
    const int64_t n_rows, n_cols; const bool own_memory; int64_t nnz = 0; const IndexBase index_base = IndexBase::Zero;
    T* vals = nullptr; sint_t* rows = nullptr; sint_t* cols = nullptr;
    NonzeroSort sort = NonzeroSort::None; bool on_device = false;
    COOMatrix(int64_t n_rows, int64_t n_cols) : n_rows(n_rows), n_cols(n_cols), own_memory(true) {}
    COOMatrix(int64_t n_rows, int64_t n_cols, int64_t nnz, T* vals, sint_t* rows, sint_t* cols)
        : n_rows(n_rows), n_cols(n_cols), own_memory(false), nnz(nnz), vals(vals), rows(rows), cols(cols) {}
    void reserve(int64_t n) {
        randblas_require(own_memory && vals == nullptr);
        nnz = n; on_device = device_pointers();
        if (n > 0) { vals = detail_sp::alloc_<T>(n, on_device); rows = detail_sp::alloc_<sint_t>(n, on_device); cols = detail_sp::alloc_<sint_t>(n, on_device); }
    }
    ~COOMatrix() { if (own_memory) { detail_sp::free_(vals, on_device); detail_sp::free_(rows, on_device); detail_sp::free_(cols, on_device); } }
    COOMatrix(const COOMatrix&) = delete;
    COOMatrix(COOMatrix&& o) : n_rows(o.n_rows), n_cols(o.n_cols), own_memory(o.own_memory), nnz(o.nnz), index_base(o.index_base), vals(o.vals),
                               rows(o.rows), cols(o.cols), sort(o.sort), on_device(o.on_device) { o.vals = nullptr; o.rows = nullptr; o.cols = nullptr; o.nnz = 0; }
};
template <typename T, typename sint_t = int64_t>
struct CSRMatrix {
    using scalar_t = T; using index_t = sint_t;
    const int64_t n_rows, n_cols; const bool own_memory; int64_t nnz = 0; const IndexBase index_base = IndexBase::Zero;
    T* vals = nullptr; sint_t* rowptr = nullptr; sint_t* colidxs = nullptr; bool on_device = false;
    CSRMatrix(int64_t n_rows, int64_t n_cols) : n_rows(n_rows), n_cols(n_cols), own_memory(true) {}
    CSRMatrix(int64_t n_rows, int64_t n_cols, int64_t nnz, T* vals, sint_t* rowptr, sint_t* colidxs)
        : n_rows(n_rows), n_cols(n_cols), own_memory(false), nnz(nnz), vals(vals), rowptr(rowptr), colidxs(colidxs) {}
    void reserve(int64_t n) {
        randblas_require(own_memory && rowptr == nullptr);
        nnz = n; on_device = device_pointers();
        rowptr = detail_sp::alloc_<sint_t>(n_rows + 1, on_device);
        if (n > 0) { vals = detail_sp::alloc_<T>(n, on_device); colidxs = detail_sp::alloc_<sint_t>(n, on_device); }
    }
    ~CSRMatrix() { if (own_memory) { detail_sp::free_(vals, on_device); detail_sp::free_(rowptr, on_device); detail_sp::free_(colidxs, on_device); } }
    CSRMatrix(const CSRMatrix&) = delete;
    CSRMatrix(CSRMatrix&& o) : n_rows(o.n_rows), n_cols(o.n_cols), own_memory(o.own_memory), nnz(o.nnz), index_base(o.index_base), vals(o.vals),
                               rowptr(o.rowptr), colidxs(o.colidxs), on_device(o.on_device) { o.vals = nullptr; o.rowptr = nullptr; o.colidxs = nullptr; o.nnz = 0; }
};
template <typename T, typename sint_t = int64_t>
struct CSCMatrix {
    using scalar_t = T; using index_t = sint_t;
    const int64_t n_rows, n_cols; const bool own_memory; int64_t nnz = 0; const IndexBase index_base = IndexBase::Zero;
    T* vals = nullptr; sint_t* rowidxs = nullptr; sint_t* colptr = nullptr; bool on_device = false;
    CSCMatrix(int64_t n_rows, int64_t n_cols) : n_rows(n_rows), n_cols(n_cols), own_memory(true) {}
    CSCMatrix(int64_t n_rows, int64_t n_cols, int64_t nnz, T* vals, sint_t* rowidxs, sint_t* colptr)
        : n_rows(n_rows), n_cols(n_cols), own_memory(false), nnz(nnz), vals(vals), rowidxs(rowidxs), colptr(colptr) {}
    void reserve(int64_t n) {
        randblas_require(own_memory && colptr == nullptr);
        nnz = n; on_device = device_pointers();
        colptr = detail_sp::alloc_<sint_t>(n_cols + 1, on_device);
        if (n > 0) { vals = detail_sp::alloc_<T>(n, on_device); rowidxs = detail_sp::alloc_<sint_t>(n, on_device); }
    }
    ~CSCMatrix() { if (own_memory) { detail_sp::free_(vals, on_device); detail_sp::free_(rowidxs, on_device); detail_sp::free_(colptr, on_device); } }
    CSCMatrix(const CSCMatrix&) = delete;
    CSCMatrix(CSCMatrix&& o) : n_rows(o.n_rows), n_cols(o.n_cols), own_memory(o.own_memory), nnz(o.nnz), index_base(o.index_base), vals(o.vals),
                               rowidxs(o.rowidxs), colptr(o.colptr), on_device(o.on_device) { o.vals = nullptr; o.rowidxs = nullptr; o.colptr = nullptr; o.nnz = 0; }
};
}  // namespace sparse_data
using sparse_data::COOMatrix; using sparse_data::CSRMatrix; using sparse_data::CSCMatrix;

namespace detail {
// elements spanned by a rows x cols matrix with leading dimension ld
inline uint64_t span(blas::Layout l, int64_t rows, int64_t cols, int64_t ld) {
    if (rows <= 0 || cols <= 0) return 0;
    return l == blas::Layout::ColMajor ? (uint64_t) ld * (cols - 1) + rows : (uint64_t) ld * (rows - 1) + cols;
}
// B as a window of a larger host array: the device copy must start from the host's contents, because the whole span is copied
// back and the reference never touches the elements between the columns / rows of the view
inline bool padded(blas::Layout l, int64_t rows, int64_t cols, int64_t ld) {
    return ld > (l == blas::Layout::ColMajor ? rows : cols) && (l == blas::Layout::ColMajor ? cols : rows) > 1;
}
template <typename sint_t>
inline std::vector<int64_t> widen(const sint_t* p, int64_t n) { return std::vector<int64_t>(p, p + n); }

#define RBX_PICK(T, name) (std::is_same<T, float>::value ? (void*) name##_f32 : (void*) name##_f64)
template <typename T> struct fn {
    using fill_dense_t = int (*)(char, int64_t, int64_t, char, char, int64_t, int64_t, int64_t, int64_t, T*, const uint32_t*, const uint32_t*, uint32_t*, void*);
    using host_fill_dense_t = int (*)(char, int64_t, int64_t, char, char, int64_t, int64_t, int64_t, int64_t, T*, const uint32_t*, const uint32_t*, uint32_t*);
    using sketch_dense_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                                   const T*, char, int64_t, int64_t, const T*, int64_t, T, T*, int64_t, void*);
    using host_sketch_dense_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                                        int64_t, int64_t, const T*, int64_t, T, T*, int64_t);
    using sketch_sparseop_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*,
                                      int64_t, int64_t, const T*, int64_t, T, T*, int64_t, void*);
    using host_sketch_sparseop_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*,
                                           int64_t, int64_t, const T*, int64_t, T, T*, int64_t);
    using spmm_coo_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, int64_t, const int64_t*, const int64_t*, const T*,
                               int64_t, int64_t, const T*, int64_t, T, T*, int64_t, void*);
    using sparsedata_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                                 int64_t, int64_t, char, int64_t, int64_t, int64_t, const T*, const int64_t*, const int64_t*, int64_t, int64_t, T, T*, int64_t, void*);
    static fill_dense_t fill_dense() { return (fill_dense_t) RBX_PICK(T, rbx_fill_dense); }
    static host_fill_dense_t host_fill_dense() { return (host_fill_dense_t) RBX_PICK(T, rbx_host_fill_dense); }
    static sketch_dense_t sketch_dense() { return (sketch_dense_t) RBX_PICK(T, rbx_sketch_dense); }
    static host_sketch_dense_t host_sketch_dense() { return (host_sketch_dense_t) RBX_PICK(T, rbx_host_sketch_dense); }
    static sketch_sparseop_t sketch_sparseop() { return (sketch_sparseop_t) RBX_PICK(T, rbx_sketch_sparseop); }
    static host_sketch_sparseop_t host_sketch_sparseop() { return (host_sketch_sparseop_t) RBX_PICK(T, rbx_host_sketch_sparseop); }
    static spmm_coo_t spmm_coo() { return (spmm_coo_t) RBX_PICK(T, rbx_spmm_coo); }
    using spmm_csx_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, int64_t, char, const int64_t*, const int64_t*, const T*,
                               const T*, int64_t, T, T*, int64_t, void*);
    static spmm_csx_t spmm_csx() { return (spmm_csx_t) RBX_PICK(T, rbx_spmm_csx); }
    static sparsedata_t sparsedata() { return (sparsedata_t) RBX_PICK(T, rbx_sketch_sparsedata); }
    using host_sparsedata_t = int (*)(char, char, char, char, int64_t, int64_t, int64_t, T, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                                      int64_t, int64_t, char, int64_t, int64_t, int64_t, const T*, const int64_t*, const int64_t*, int64_t, int64_t, T, T*, int64_t);
    static host_sparsedata_t host_sparsedata() { return (host_sparsedata_t) RBX_PICK(T, rbx_host_sketch_sparsedata); }
};
#undef RBX_PICK
}  // namespace detail

// ---------------------------------------------------------------------------------------------------
// fill_dense (dense_skops.hh:486-592) and fill_sparse (sparse_skops.hh:389-413)
// ---------------------------------------------------------------------------------------------------
template <typename T, typename RNG = r123::Philox4x32>
inline RNGState<RNG> fill_dense(blas::Layout layout, const DenseDist& D, int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s,
                                T* buff, const RNGState<RNG>& seed) {
    static_assert(detail::is_real<T>::value, "T must be float or double");
    if (D.family == DenseDistName::BlackBox) throw std::invalid_argument("fill_dense cannot be called with the BlackBox distribution.");
    RNGState<RNG> next = seed;
    const int rc = device_pointers()
        ? detail::fn<T>::fill_dense()(detail::ch(layout), D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis, n_rows, n_cols, ro_s, co_s,
                                      buff, seed.counter.v, seed.key.v, next.counter.v, current_stream())
        : detail::fn<T>::host_fill_dense()(detail::ch(layout), D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis, n_rows, n_cols, ro_s, co_s,
                                           buff, seed.counter.v, seed.key.v, next.counter.v);
    detail::check(rc, "fill_dense");
    return next;
}
template <typename T, typename RNG = r123::Philox4x32>
inline RNGState<RNG> fill_dense(const DenseDist& D, T* buff, const RNGState<RNG>& seed) {
    return fill_dense(dist_to_layout(D), D, D.n_rows, D.n_cols, 0, 0, buff, seed);
}
template <typename DenseSkOpT>
inline void fill_dense(DenseSkOpT& S) {
    using T = typename DenseSkOpT::buffer_type;
    randblas_require(S.buff == nullptr);
    const uint64_t n = (uint64_t) S.dist.n_rows * S.dist.n_cols;
    if (device_pointers()) { void* p = nullptr; detail::check(rbx_malloc(&p, n * sizeof(T)), "rbx_malloc"); S.buff = (T*) p; }
    else S.buff = new T[n];
    S.del_buff_on_destruct = true;
    fill_dense(S.dist, S.buff, S.seed_state);
}

template <typename SparseSkOpT>
inline void fill_sparse(SparseSkOpT& S) {
    using T = typename SparseSkOpT::scalar_t; using I = typename SparseSkOpT::index_t;
    const int64_t n = nnz(S);
    uint32_t next[4];
    auto call = [&](I* r, I* c, T* v) -> int {
        if constexpr (std::is_same<T, float>::value && std::is_same<I, int32_t>::value)
            return rbx_fill_sparse_f32_i32(S.dist.n_rows, S.dist.n_cols, S.dist.vec_nnz, (char) S.dist.major_axis, S.seed_state.counter.v, S.seed_state.key.v, r, c, v, next, current_stream());
        else if constexpr (std::is_same<T, float>::value)
            return rbx_fill_sparse_f32_i64(S.dist.n_rows, S.dist.n_cols, S.dist.vec_nnz, (char) S.dist.major_axis, S.seed_state.counter.v, S.seed_state.key.v, r, c, v, next, current_stream());
        else if constexpr (std::is_same<I, int32_t>::value)
            return rbx_fill_sparse_f64_i32(S.dist.n_rows, S.dist.n_cols, S.dist.vec_nnz, (char) S.dist.major_axis, S.seed_state.counter.v, S.seed_state.key.v, r, c, v, next, current_stream());
        else
            return rbx_fill_sparse_f64_i64(S.dist.n_rows, S.dist.n_cols, S.dist.vec_nnz, (char) S.dist.major_axis, S.seed_state.counter.v, S.seed_state.key.v, r, c, v, next, current_stream());
    };
    if (device_pointers()) {
        detail::check(call(S.rows, S.cols, S.vals), "fill_sparse");
    } else {
        detail::DevBuf r, c, v;
        r.alloc(n * sizeof(I)); c.alloc(n * sizeof(I)); v.alloc(n * sizeof(T));
        detail::check(call((I*) r.p, (I*) c.p, (T*) v.p), "fill_sparse");
        detail::check(rbx_memcpy_d2h(S.rows, r.p, n * sizeof(I), nullptr), "rbx_memcpy_d2h");
        detail::check(rbx_memcpy_d2h(S.cols, c.p, n * sizeof(I), nullptr), "rbx_memcpy_d2h");
        detail::check(rbx_memcpy_d2h(S.vals, v.p, n * sizeof(T), nullptr), "rbx_memcpy_d2h");
        detail::check(rbx_stream_synchronize(nullptr), "rbx_stream_synchronize");
    }
    S.known_filled = true;
}

// ---------------------------------------------------------------------------------------------------
// cores: dense::lskge3 / rskge3 (skge.hh:173-215, 320-364), sparse::lskges / rskges (skge.hh:485-510, 616-641)
// side 'L': dims (d, n, m); side 'R': dims (m, d, n)   -- see rbx.h
// ---------------------------------------------------------------------------------------------------
namespace detail {
inline void ab_dims(char side, blas::Op opA, int64_t p0, int64_t p1, int64_t p2, int64_t& rA, int64_t& cA, int64_t& rB, int64_t& cB) {
    if (side == 'L') { rA = (opA == blas::Op::NoTrans) ? p2 : p1; cA = (opA == blas::Op::NoTrans) ? p1 : p2; rB = p0; cB = p1; }
    else             { rA = (opA == blas::Op::NoTrans) ? p0 : p2; cA = (opA == blas::Op::NoTrans) ? p2 : p0; rB = p0; cB = p1; }
}

template <typename T, typename RNG>
inline void sk3(char side, blas::Layout layout, blas::Op opS, blas::Op opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                DenseSkOp<T, RNG>& S, int64_t ro_s, int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb) {
    const DenseDist& D = S.dist;
    if (device_pointers()) {
        check(fn<T>::sketch_dense()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis,
                                    S.seed_state.counter.v, S.seed_state.key.v, S.buff, ch(S.layout), ro_s, co_s, A, lda, beta, B, ldb, current_stream()),
              "sketch_general");
    } else if (S.buff == nullptr) {
        check(fn<T>::host_sketch_dense()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis,
                                         S.seed_state.counter.v, S.seed_state.key.v, ro_s, co_s, A, lda, beta, B, ldb),
              "sketch_general");
    } else {
        // explicit host operator (DenseSkOp::buff, possibly BlackBox): validate first (null data pointers => dry run is not
        // available here, so let the device call validate), staging S, A and B by hand
        int64_t rA, cA, rB, cB; ab_dims(side, opA, p0, p1, p2, rA, cA, rB, cB);
        randblas_require(lda >= (layout == blas::Layout::ColMajor ? rA : cA));
        randblas_require(ldb >= (layout == blas::Layout::ColMajor ? rB : cB));
        const uint64_t nA = span(layout, rA, cA, lda), nB = span(layout, rB, cB, ldb);
        DevBuf dS(S.buff, (uint64_t) D.n_rows * D.n_cols * sizeof(T)), dA(A, nA * sizeof(T)),
               dB((beta != (T) 0 || padded(layout, rB, cB, ldb)) ? B : nullptr, nB * sizeof(T));
        check(fn<T>::sketch_dense()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis,
                                    S.seed_state.counter.v, S.seed_state.key.v, (const T*) dS.p, ch(S.layout), ro_s, co_s, (const T*) dA.p, lda, beta,
                                    (T*) dB.p, ldb, nullptr), "sketch_general");
        check(rbx_memcpy_d2h(B, dB.p, nB * sizeof(T), nullptr), "rbx_memcpy_d2h");
        check(rbx_stream_synchronize(nullptr), "rbx_stream_synchronize");
    }
}

template <typename T, typename RNG, typename sint_t>
inline void sk3(char side, blas::Layout layout, blas::Op opS, blas::Op opA, int64_t p0, int64_t p1, int64_t p2, T alpha,
                SparseSkOp<T, RNG, sint_t>& S, int64_t ro_s, int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb) {
    const SparseDist& D = S.dist;
    if (!S.known_filled) {
        // the reference fills the operator here (skge.hh:503-504); on the device it is regenerated inside the sketch
        const int rc = device_pointers()
            ? fn<T>::sketch_sparseop()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, D.vec_nnz, (char) D.major_axis,
                                       S.seed_state.counter.v, S.seed_state.key.v, ro_s, co_s, A, lda, beta, B, ldb, current_stream())
            : fn<T>::host_sketch_sparseop()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, D.vec_nnz, (char) D.major_axis,
                                            S.seed_state.counter.v, S.seed_state.key.v, ro_s, co_s, A, lda, beta, B, ldb);
        check(rc, "sketch_general");
        return;
    }
    // explicit triplets (the caller may have edited them): COO spmm, spmm_dispatch.hh:48-200
    const int64_t n = nnz(S);
    if (device_pointers()) {
        if constexpr (std::is_same<sint_t, int64_t>::value) {
            check(fn<T>::spmm_coo()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, n, S.rows, S.cols, S.vals,
                                    ro_s, co_s, A, lda, beta, B, ldb, current_stream()), "sketch_general");
        } else {
            throw std::runtime_error("explicit 32-bit COO operators need 64-bit indices on the device; use an unfilled SparseSkOp");
        }
        return;
    }
    int64_t rA, cA, rB, cB; ab_dims(side, opA, p0, p1, p2, rA, cA, rB, cB);
    randblas_require(lda >= (layout == blas::Layout::ColMajor ? rA : cA));
    randblas_require(ldb >= (layout == blas::Layout::ColMajor ? rB : cB));
    const uint64_t nA = span(layout, rA, cA, lda), nB = span(layout, rB, cB, ldb);
    const std::vector<int64_t> r64 = widen(S.rows, n), c64 = widen(S.cols, n);
    DevBuf dr(r64.data(), n * 8), dc(c64.data(), n * 8), dv(S.vals, n * sizeof(T)), dA(A, nA * sizeof(T)),
           dB((beta != (T) 0 || padded(layout, rB, cB, ldb)) ? B : nullptr, nB * sizeof(T));
    check(fn<T>::spmm_coo()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, n, (const int64_t*) dr.p, (const int64_t*) dc.p,
                            (const T*) dv.p, ro_s, co_s, (const T*) dA.p, lda, beta, (T*) dB.p, ldb, nullptr), "sketch_general");
    check(rbx_memcpy_d2h(B, dB.p, nB * sizeof(T), nullptr), "rbx_memcpy_d2h");
    check(rbx_stream_synchronize(nullptr), "rbx_stream_synchronize");
}
}  // namespace detail

namespace dense {
template <typename T, typename SKOP>
inline void lskge3(blas::Layout layout, blas::Op opS, blas::Op opA, int64_t d, int64_t n, int64_t m, T alpha, SKOP& S, int64_t ro_s, int64_t co_s,
                   const T* A, int64_t lda, T beta, T* B, int64_t ldb) { detail::sk3('L', layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); }
template <typename T, typename SKOP>
inline void rskge3(blas::Layout layout, blas::Op opA, blas::Op opS, int64_t m, int64_t d, int64_t n, T alpha, const T* A, int64_t lda, SKOP& S,
                   int64_t ro_s, int64_t co_s, T beta, T* B, int64_t ldb) { detail::sk3('R', layout, opS, opA, m, d, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); }
}  // namespace dense
namespace sparse {
template <typename T, typename SKOP>
inline void lskges(blas::Layout layout, blas::Op opS, blas::Op opA, int64_t d, int64_t n, int64_t m, T alpha, SKOP& S, int64_t ro_s, int64_t co_s,
                   const T* A, int64_t lda, T beta, T* B, int64_t ldb) { detail::sk3('L', layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); }
template <typename T, typename SKOP>
inline void rskges(blas::Layout layout, blas::Op opA, blas::Op opS, int64_t m, int64_t d, int64_t n, T alpha, const T* A, int64_t lda, SKOP& S,
                   int64_t ro_s, int64_t co_s, T beta, T* B, int64_t ldb) { detail::sk3('R', layout, opS, opA, m, d, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); }
}  // namespace sparse

// ---------------------------------------------------------------------------------------------------
// sketch_general (skge.hh:771-1214): left / right, with and without sub-matrix offsets; SKOP = DenseSkOp or SparseSkOp
// ---------------------------------------------------------------------------------------------------
template <typename T, typename SKOP>
inline void sketch_general(blas::Layout layout, blas::Op opS, blas::Op opA, int64_t d, int64_t n, int64_t m, T alpha, SKOP& S, int64_t ro_s,
                           int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb) {
    detail::sk3('L', layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, lda, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_general(blas::Layout layout, blas::Op opA, blas::Op opS, int64_t m, int64_t d, int64_t n, T alpha, const T* A, int64_t lda,
                           SKOP& S, int64_t ro_s, int64_t co_s, T beta, T* B, int64_t ldb) {
    detail::sk3('R', layout, opS, opA, m, d, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_general(blas::Layout layout, blas::Op opS, blas::Op opA, int64_t d, int64_t n, int64_t m, T alpha, SKOP& S, const T* A,
                           int64_t lda, T beta, T* B, int64_t ldb) {
    if (opS == blas::Op::NoTrans) { randblas_require(S.dist.n_rows == d); randblas_require(S.dist.n_cols == m); }
    else                          { randblas_require(S.dist.n_rows == m); randblas_require(S.dist.n_cols == d); }      // skge.hh:1104-1111
    sketch_general(layout, opS, opA, d, n, m, alpha, S, (int64_t) 0, (int64_t) 0, A, lda, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_general(blas::Layout layout, blas::Op opA, blas::Op opS, int64_t m, int64_t d, int64_t n, T alpha, const T* A, int64_t lda,
                           SKOP& S, T beta, T* B, int64_t ldb) {
    if (opS == blas::Op::NoTrans) { randblas_require(S.dist.n_rows == n); randblas_require(S.dist.n_cols == d); }
    else                          { randblas_require(S.dist.n_rows == d); randblas_require(S.dist.n_cols == n); }      // skge.hh:1206-1213
    sketch_general(layout, opA, opS, m, d, n, alpha, A, lda, S, (int64_t) 0, (int64_t) 0, beta, B, ldb);
}

// ---------------------------------------------------------------------------------------------------
// sketch_vector (skve.hh:152-258)
// ---------------------------------------------------------------------------------------------------
template <typename T, typename SKOP>
inline void sketch_vector(blas::Op opS, int64_t d, int64_t m, T alpha, SKOP& S, int64_t ro_s, int64_t co_s, const T* x, int64_t incx, T beta,
                          T* y, int64_t incy) {
    int64_t _d = d, _m = m;
    if (opS == blas::Op::Trans) { _d = m; _m = d; }
    sketch_general(blas::Layout::RowMajor, opS, blas::Op::NoTrans, _d, (int64_t) 1, _m, alpha, S, ro_s, co_s, x, incx, beta, y, incy);
}
template <typename T, typename SKOP>
inline void sketch_vector(blas::Op opS, T alpha, SKOP& S, const T* x, int64_t incx, T beta, T* y, int64_t incy) {
    sketch_vector(opS, S.dist.n_rows, S.dist.n_cols, alpha, S, (int64_t) 0, (int64_t) 0, x, incx, beta, y, incy);
}

// ---------------------------------------------------------------------------------------------------
// sketch_symmetric (sksy.hh:165-537); util::require_symmetric (util.hh:165-188) runs on the host copy of A
// ---------------------------------------------------------------------------------------------------
namespace util {
template <typename T>
inline void require_symmetric(blas::Layout layout, const T* A, int64_t n, int64_t lda, T tol) {
    if (tol < 0) return;
    if (device_pointers()) {                       // device-resident A: one kernel over the upper triangle
        const int rc = std::is_same<T, float>::value
            ? rbx_require_symmetric_f32(detail::ch(layout), (const float*) A, n, lda, (float) tol, current_stream())
            : rbx_require_symmetric_f64(detail::ch(layout), (const double*) A, n, lda, (double) tol, current_stream());
        if (rc == RBX_ERR_ARG) throw Error("Symmetry check failed, in function require_symmetric");
        detail::check(rc, "require_symmetric");
        return;
    }
    const int64_t rs = layout == blas::Layout::ColMajor ? 1 : lda, cs = layout == blas::Layout::ColMajor ? lda : 1;
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = i + 1; j < n; ++j) {
            const T Aij = A[i * rs + j * cs], Aji = A[j * rs + i * cs];
            const T viol = std::abs(Aij - Aji), rel_tol = (std::abs(Aij) + std::abs(Aji) + 1) * tol;
            if (viol > rel_tol) throw Error("Symmetry check failed, in function require_symmetric");
        }
}
// ---- index sampling (util.hh:257-334): same Philox + uneg11 stream and returned state as the reference ----------
// run body(dev_ptr) on a device copy of a host array of `bytes` bytes (in/out), or on the pointer itself in device mode
template <typename F>
inline void with_device_(void* host_or_dev, uint64_t bytes, bool copy_in, bool copy_out, F body) {
    if (device_pointers() || bytes == 0) { body(host_or_dev); return; }
    detail::DevBuf d;
    if (copy_in) d.upload(host_or_dev, bytes); else d.alloc(bytes);
    body(d.p);
    if (copy_out) detail::check(rbx_memcpy_d2h(host_or_dev, d.p, bytes, nullptr), "rbx_memcpy_d2h");
}

template <typename T>
inline void weights_to_cdf(int64_t n, T* w, T error_if_below = -std::numeric_limits<T>::epsilon()) {      // util.hh:257-270
    static_assert(detail::is_real<T>::value, "float or double");
    with_device_(w, (uint64_t) n * sizeof(T), true, true, [&](void* d) {
        if (std::is_same<T, float>::value) detail::check(rbx_weights_to_cdf_f32(n, (float*) d, (float) error_if_below, current_stream()), "weights_to_cdf");
        else                               detail::check(rbx_weights_to_cdf_f64(n, (double*) d, (double) error_if_below, current_stream()), "weights_to_cdf");
    });
}
template <typename TF, typename RNG = r123::Philox4x32>
inline RNGState<RNG> sample_indices_iid(int64_t n, TF* cdf, int64_t k, int64_t* samples, RNGState<RNG> state) {   // util.hh:284-305
    static_assert(detail::is_real<TF>::value, "float or double");
    RNGState<RNG> out = state;
    with_device_(cdf, (uint64_t) n * sizeof(TF), true, false, [&](void* dc) {
        with_device_(samples, (uint64_t) k * 8, false, true, [&](void* ds) {
            if (std::is_same<TF, float>::value)
                detail::check(rbx_sample_indices_iid_f32(n, (const float*) dc, k, (int64_t*) ds, state.counter.v, state.key.v, out.counter.v, current_stream()), "sample_indices_iid");
            else
                detail::check(rbx_sample_indices_iid_f64(n, (const double*) dc, k, (int64_t*) ds, state.counter.v, state.key.v, out.counter.v, current_stream()), "sample_indices_iid");
        });
    });
    return out;
}
template <typename RNG = r123::Philox4x32>
inline RNGState<RNG> sample_indices_iid_uniform(int64_t n, int64_t k, int64_t* samples, RNGState<RNG> state) {     // util.hh:311-334
    RNGState<RNG> out = state;
    with_device_(samples, (uint64_t) k * 8, false, true, [&](void* ds) {
        detail::check(rbx_sample_indices_iid_uniform(n, k, (int64_t*) ds, state.counter.v, state.key.v, out.counter.v, current_stream()), "sample_indices_iid_uniform");
    });
    return out;
}
}  // namespace util

template <typename T, typename SKOP>
inline void sketch_symmetric(blas::Layout layout, int64_t n, int64_t d, T alpha, const T* A, int64_t lda, SKOP& S, int64_t ro_s, int64_t co_s,
                             T beta, T* B, int64_t ldb, T sym_check_tol = 0) {
    util::require_symmetric(layout, A, n, lda, sym_check_tol);
    sketch_general(layout, blas::Op::NoTrans, blas::Op::NoTrans, n, d, n, alpha, A, lda, S, ro_s, co_s, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_symmetric(blas::Layout layout, int64_t d, int64_t n, T alpha, SKOP& S, int64_t ro_s, int64_t co_s, const T* A, int64_t lda,
                             T beta, T* B, int64_t ldb, T sym_check_tol = 0) {
    util::require_symmetric(layout, A, n, lda, sym_check_tol);
    sketch_general(layout, blas::Op::NoTrans, blas::Op::NoTrans, d, n, n, alpha, S, ro_s, co_s, A, lda, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_symmetric(blas::Layout layout, T alpha, const T* A, int64_t lda, SKOP& S, T beta, T* B, int64_t ldb, T sym_check_tol = 0) {
    const int64_t n = S.dist.n_rows, d = S.dist.n_cols;
    util::require_symmetric(layout, A, n, lda, sym_check_tol);
    sketch_general(layout, blas::Op::NoTrans, blas::Op::NoTrans, n, d, n, alpha, A, lda, S, (int64_t) 0, (int64_t) 0, beta, B, ldb);
}
template <typename T, typename SKOP>
inline void sketch_symmetric(blas::Layout layout, T alpha, SKOP& S, const T* A, int64_t lda, T beta, T* B, int64_t ldb, T sym_check_tol = 0) {
    const int64_t d = S.dist.n_rows, n = S.dist.n_cols;
    util::require_symmetric(layout, A, n, lda, sym_check_tol);
    sketch_general(layout, blas::Op::NoTrans, blas::Op::NoTrans, d, n, n, alpha, S, (int64_t) 0, (int64_t) 0, A, lda, beta, B, ldb);
}

// ---------------------------------------------------------------------------------------------------
// sketch_sparse (sksp.hh:464-616): dense operator applied to sparse data (COO / CSR / CSC)
// ---------------------------------------------------------------------------------------------------
namespace detail {
template <typename T, typename I> inline void sp_arrays(const COOMatrix<T, I>& A, char& fmt, const I*& i0, int64_t& n0, const I*& i1, int64_t& n1) {
    fmt = 'O'; i0 = A.rows; n0 = A.nnz; i1 = A.cols; n1 = A.nnz; }
template <typename T, typename I> inline void sp_arrays(const CSRMatrix<T, I>& A, char& fmt, const I*& i0, int64_t& n0, const I*& i1, int64_t& n1) {
    fmt = 'R'; i0 = A.rowptr; n0 = A.n_rows + 1; i1 = A.colidxs; n1 = A.nnz; }
template <typename T, typename I> inline void sp_arrays(const CSCMatrix<T, I>& A, char& fmt, const I*& i0, int64_t& n0, const I*& i1, int64_t& n1) {
    fmt = 'C'; i0 = A.rowidxs; n0 = A.nnz; i1 = A.colptr; n1 = A.n_cols + 1; }

template <typename T, typename RNG, typename SpMat>
inline void sksp3(char side, blas::Layout layout, blas::Op opS, blas::Op opA, int64_t p0, int64_t p1, int64_t p2, T alpha, DenseSkOp<T, RNG>& S,
                  int64_t ro_s, int64_t co_s, SpMat& A, int64_t ro_a, int64_t co_a, T beta, T* B, int64_t ldb) {
    using I = typename SpMat::index_t;
    randblas_require(A.index_base == sparse_data::IndexBase::Zero);           // spmm_dispatch.hh:89
    randblas_require(S.buff == nullptr);                                       // the device path regenerates S from its seed
    const DenseDist& D = S.dist;
    char fmt; const I *i0, *i1; int64_t n0, n1;
    sp_arrays(A, fmt, i0, n0, i1, n1);
    if (device_pointers()) {
        if constexpr (std::is_same<I, int64_t>::value) {
            check(fn<T>::sparsedata()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis,
                                      S.seed_state.counter.v, S.seed_state.key.v, ro_s, co_s, fmt, A.n_rows, A.n_cols, A.nnz, A.vals, i0, i1, ro_a, co_a,
                                      beta, B, ldb, current_stream()), "sketch_sparse");
        } else {
            throw std::runtime_error("sketch_sparse on device pointers needs 64-bit indices");
        }
        return;
    }
    // host arrays: the library stages the matrix and B (rbx_host_sketch_sparsedata_*); 32-bit indices are widened first
    const std::vector<int64_t> a0 = widen(i0, n0), a1 = widen(i1, n1);
    check(fn<T>::host_sparsedata()(side, ch(layout), ch(opS), ch(opA), p0, p1, p2, alpha, D.n_rows, D.n_cols, (char) D.family, (char) D.major_axis,
                                   S.seed_state.counter.v, S.seed_state.key.v, ro_s, co_s, fmt, A.n_rows, A.n_cols, A.nnz, A.vals,
                                   a0.data(), a1.data(), ro_a, co_a, beta, B, ldb), "sketch_sparse");
}
}  // namespace detail

template <typename T, typename RNG, typename SpMat>
inline void sketch_sparse(blas::Layout layout, blas::Op opS, blas::Op opA, int64_t d, int64_t n, int64_t m, T alpha, DenseSkOp<T, RNG>& S,
                          int64_t ro_s, int64_t co_s, SpMat& A, int64_t ro_a, int64_t co_a, T beta, T* B, int64_t ldb) {
    detail::sksp3('L', layout, opS, opA, d, n, m, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb);
}
template <typename T, typename RNG, typename SpMat>
inline void sketch_sparse(blas::Layout layout, blas::Op opA, blas::Op opS, int64_t m, int64_t d, int64_t n, T alpha, SpMat& A, int64_t ro_a,
                          int64_t co_a, DenseSkOp<T, RNG>& S, int64_t ro_s, int64_t co_s, T beta, T* B, int64_t ldb) {
    detail::sksp3('R', layout, opS, opA, m, d, n, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb);
}


// ---------------------------------------------------------------------------------------------------
// sparse_skops.hh:484-513: COO view of a sparse operator, and its transpose (both alias S.rows / S.cols / S.vals)
// ---------------------------------------------------------------------------------------------------
namespace sparse {
template <typename SkOp, typename T = typename SkOp::scalar_t, typename sint_t = typename SkOp::index_t>
inline COOMatrix<T, sint_t> coo_view_of_skop(SkOp& S) {
    if (!S.known_filled) fill_sparse(S);
    return COOMatrix<T, sint_t>(S.dist.n_rows, S.dist.n_cols, nnz(S), S.vals, S.rows, S.cols);
}
template <typename SKOP>
inline SKOP transpose(SKOP const& S) {
    randblas_require(S.known_filled);
    SparseDist dist = { S.dist.n_cols, S.dist.n_rows, S.dist.vec_nnz, S.dist.major_axis };
    SKOP St(dist, S.seed_state, S.cols, S.rows, S.vals, S.next_state);
    St.known_filled = true;
    return St;
}
}  // namespace sparse
using sparse::coo_view_of_skop;
using sparse::transpose;

// ---------------------------------------------------------------------------------------------------
// sparse_data/spmm_dispatch.hh:48-200, 290-385: left_spmm / right_spmm / spmm -- sparse (COO / CSR / CSC) times dense.
// The product runs on the device as an output-stationary gather (rbx_spmm_coo_*: no atomics, fixed summation order).
// ---------------------------------------------------------------------------------------------------
namespace sparse_data {
namespace detail_spmm {
// COO triplets (int64, device) of a sparse matrix in any of the three formats; owns whatever it had to create
template <typename T>
struct DevCoo {
    const int64_t* rows = nullptr; const int64_t* cols = nullptr; const T* vals = nullptr; int64_t nnz = 0;
    ::RandBLAS::detail::DevBuf b0, b1, bv, bx;
};
template <typename T, typename I>
inline void outer_indices(int64_t n_outer, const I* ptr_host, std::vector<int64_t>& out) {
    out.resize((size_t) ptr_host[n_outer]);
    for (int64_t o = 0; o < n_outer; ++o) for (I e = ptr_host[o]; e < ptr_host[o + 1]; ++e) out[(size_t) e] = o;
}
template <typename T, typename SpMat>
inline void to_dev_coo(SpMat& A, DevCoo<T>& d) {
    using I = typename SpMat::index_t;
    char fmt; const I *i0, *i1; int64_t n0, n1;
    ::RandBLAS::detail::sp_arrays(A, fmt, i0, n0, i1, n1);
    d.nnz = A.nnz;
    using ::RandBLAS::detail::check;
    if (device_pointers()) {
        if constexpr (!std::is_same<I, int64_t>::value) { throw std::runtime_error("sparse matrices on device pointers need 64-bit indices"); }
        else {
            d.vals = A.vals;
            if (fmt == 'O') { d.rows = i0; d.cols = i1; return; }
            d.bx.alloc((uint64_t) (A.nnz > 0 ? A.nnz : 1) * 8);
            const int64_t n_outer = fmt == 'R' ? A.n_rows : A.n_cols;
            check(rbx_expand_ptr(n_outer, fmt == 'R' ? i0 : i1, (int64_t*) d.bx.p, current_stream()), "rbx_expand_ptr");
            if (fmt == 'R') { d.rows = (const int64_t*) d.bx.p; d.cols = i1; } else { d.rows = i0; d.cols = (const int64_t*) d.bx.p; }
        }
        return;
    }
    std::vector<int64_t> r, c;
    if (fmt == 'O') { r.assign(i0, i0 + A.nnz); c.assign(i1, i1 + A.nnz); }
    else if (fmt == 'R') { outer_indices<T>(A.n_rows, i0, r); c.assign(i1, i1 + A.nnz); }
    else { r.assign(i0, i0 + A.nnz); outer_indices<T>(A.n_cols, i1, c); }
    const uint64_t n = (uint64_t) (A.nnz > 0 ? A.nnz : 1);
    d.b0.upload(A.nnz ? r.data() : nullptr, n * 8); d.b1.upload(A.nnz ? c.data() : nullptr, n * 8); d.bv.upload(A.nnz ? A.vals : nullptr, n * sizeof(T));
    d.rows = (const int64_t*) d.b0.p; d.cols = (const int64_t*) d.b1.p; d.vals = (const T*) d.bv.p;
}
}  // namespace detail_spmm

template <typename T, typename SpMat>
inline void left_spmm(blas::Layout layout, blas::Op opA, blas::Op opB, int64_t d, int64_t n, int64_t m, T alpha, SpMat& A, int64_t ro_a, int64_t co_a,
                      const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
    using namespace ::RandBLAS::detail;
    randblas_require(A.index_base == IndexBase::Zero);                                        // spmm_dispatch.hh:89
    const int64_t rows_sub = opA == blas::Op::NoTrans ? d : m, cols_sub = opA == blas::Op::NoTrans ? m : d;
    char fmt; const typename SpMat::index_t *i0, *i1; int64_t n0, n1;
    sp_arrays(A, fmt, i0, n0, i1, n1);
    if (fmt == 'O') { randblas_require(A.n_rows >= rows_sub + ro_a); randblas_require(A.n_cols >= cols_sub + co_a); }     // :94-96
    else { randblas_require(A.n_rows == rows_sub); randblas_require(A.n_cols == cols_sub); randblas_require(ro_a == 0); randblas_require(co_a == 0); }  // :97-104
    if constexpr (std::is_same<typename SpMat::index_t, int64_t>::value) {
        if (fmt != 'O' && device_pointers()) {
            // device-resident CSR / CSC: the arrays go down as they are (used in place when compressed along the output axis)
            check(::RandBLAS::detail::fn<T>::spmm_csx()('L', ch(layout), ch(opA), ch(opB), d, n, m, alpha, A.n_rows, A.n_cols, A.nnz, fmt,
                                                        fmt == 'R' ? i0 : i1, fmt == 'R' ? i1 : i0, A.vals, B, ldb, beta, C, ldc, current_stream()), "left_spmm");
            return;
        }
    }
    detail_spmm::DevCoo<T> dc;
    detail_spmm::to_dev_coo<T>(A, dc);
    auto call = ::RandBLAS::detail::fn<T>::spmm_coo();
    if (device_pointers()) {
        check(call('L', ch(layout), ch(opA), ch(opB), d, n, m, alpha, A.n_rows, A.n_cols, dc.nnz, dc.rows, dc.cols, dc.vals, ro_a, co_a, B, ldb, beta, C, ldc,
                 current_stream()), "left_spmm");
        if (dc.bx.p) check(rbx_stream_synchronize(current_stream()), "rbx_stream_synchronize");     // the expanded indices are freed on return
        return;
    }
    const int64_t rB = opB == blas::Op::NoTrans ? m : n, cB = opB == blas::Op::NoTrans ? n : m;
    randblas_require(ldb >= (layout == blas::Layout::ColMajor ? rB : cB));
    randblas_require(ldc >= (layout == blas::Layout::ColMajor ? d : n));
    const uint64_t nB = span(layout, rB, cB, ldb), nC = span(layout, d, n, ldc);
    DevBuf dB(B, nB * sizeof(T)), dC((beta != (T) 0 || padded(layout, d, n, ldc)) ? C : nullptr, nC * sizeof(T));
    check(call('L', ch(layout), ch(opA), ch(opB), d, n, m, alpha, A.n_rows, A.n_cols, dc.nnz, dc.rows, dc.cols, dc.vals, ro_a, co_a, (const T*) dB.p, ldb, beta,
             (T*) dC.p, ldc, nullptr), "left_spmm");
    check(rbx_memcpy_d2h(C, dC.p, nC * sizeof(T), nullptr), "rbx_memcpy_d2h");
    check(rbx_stream_synchronize(nullptr), "rbx_stream_synchronize");
}

// C (m x d) = alpha * op(A) (m x n, dense) * op(submat(B)) (n x d, sparse) + beta * C: the transposed left product in the
// opposite layout, exactly as spmm_dispatch.hh:162-200 reduces it
template <typename T, typename SpMat>
inline void right_spmm(blas::Layout layout, blas::Op opA, blas::Op opB, int64_t m, int64_t d, int64_t n, T alpha, const T* A, int64_t lda, SpMat& B,
                       int64_t i_off, int64_t j_off, T beta, T* C, int64_t ldc) {
    const blas::Op trans_opB = opB == blas::Op::NoTrans ? blas::Op::Trans : blas::Op::NoTrans;
    const blas::Layout trans_layout = layout == blas::Layout::ColMajor ? blas::Layout::RowMajor : blas::Layout::ColMajor;
    left_spmm(trans_layout, trans_opB, opA, d, m, n, alpha, B, i_off, j_off, A, lda, beta, C, ldc);
}
template <typename T, typename SpMat, typename std::enable_if<!std::is_pointer<SpMat>::value && !std::is_arithmetic<SpMat>::value, int>::type = 0>
inline void spmm(blas::Layout layout, blas::Op opA, blas::Op opB, int64_t m, int64_t n, int64_t k, T alpha, SpMat& A, int64_t ro_a, int64_t co_a,
                 const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
    left_spmm(layout, opA, opB, m, n, k, alpha, A, ro_a, co_a, B, ldb, beta, C, ldc);
}
template <typename T, typename SpMat>
inline void spmm(blas::Layout layout, blas::Op opA, blas::Op opB, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda, SpMat& B,
                 int64_t ro_b, int64_t co_b, T beta, T* C, int64_t ldc) {
    right_spmm(layout, opA, opB, m, n, k, alpha, A, lda, B, ro_b, co_b, beta, C, ldc);     // (the reference's overload passes a stray argument, spmm_dispatch.hh:382)
}

// ---------------------------------------------------------------------------------------------------
// sparse_data/conversions.hh:43-159. The sort and the compression run on the device (rbx_sparse_regroup_*); inner
// indices come out increasing in every row / column, so the result does not depend on the order of the input triplets.
// ---------------------------------------------------------------------------------------------------
namespace conversions {
namespace detail_cv {
template <typename T, typename SpIn>
inline void regroup(char by, SpIn& in, int64_t n_outer, int64_t* ptr_out, int64_t* idx_out, T* vals_out, bool out_on_device) {
    using ::RandBLAS::detail::check; using ::RandBLAS::detail::DevBuf;
    detail_spmm::DevCoo<T> dc;
    detail_spmm::to_dev_coo<T>(in, dc);
    auto call = [&](int64_t* p, int64_t* i, T* v, void* st) {
        if (std::is_same<T, float>::value) check(rbx_sparse_regroup_f32(by, 'O', in.n_rows, in.n_cols, dc.nnz, (const float*) dc.vals, dc.rows, dc.cols, p, i, (float*) v, st), "regroup");
        else check(rbx_sparse_regroup_f64(by, 'O', in.n_rows, in.n_cols, dc.nnz, (const double*) dc.vals, dc.rows, dc.cols, p, i, (double*) v, st), "regroup");
    };
    if (out_on_device) { call(ptr_out, idx_out, vals_out, current_stream()); check(rbx_stream_synchronize(current_stream()), "rbx_stream_synchronize"); return; }
    const uint64_t n = (uint64_t) (dc.nnz > 0 ? dc.nnz : 1);
    DevBuf dp, di, dv;
    dp.alloc((uint64_t) (n_outer + 1) * 8); di.alloc(n * 8); dv.alloc(n * sizeof(T));
    call((int64_t*) dp.p, (int64_t*) di.p, (T*) dv.p, nullptr);
    check(rbx_memcpy_d2h(ptr_out, dp.p, (uint64_t) (n_outer + 1) * 8, nullptr), "rbx_memcpy_d2h");
    if (dc.nnz > 0) { check(rbx_memcpy_d2h(idx_out, di.p, n * 8, nullptr), "rbx_memcpy_d2h"); check(rbx_memcpy_d2h(vals_out, dv.p, n * sizeof(T), nullptr), "rbx_memcpy_d2h"); }
    check(rbx_stream_synchronize(nullptr), "rbx_stream_synchronize");
}
template <typename I> inline void narrow(const std::vector<int64_t>& src, I* dst) { for (size_t i = 0; i < src.size(); ++i) dst[i] = (I) src[i]; }
}  // namespace detail_cv

template <typename T, typename sint_t1 = int64_t, typename sint_t2 = int64_t>
inline void coo_to_csc(COOMatrix<T, sint_t1>& coo, CSCMatrix<T, sint_t2>& csc) {
    randblas_require(csc.n_rows == coo.n_rows); randblas_require(csc.n_cols == coo.n_cols);
    randblas_require(csc.index_base == IndexBase::Zero); randblas_require(coo.index_base == IndexBase::Zero);
    csc.reserve(coo.nnz);
    if (std::is_same<sint_t2, int64_t>::value) { detail_cv::regroup<T>('C', coo, coo.n_cols, (int64_t*) csc.colptr, (int64_t*) csc.rowidxs, csc.vals, csc.on_device); return; }
    randblas_require(!csc.on_device);
    std::vector<int64_t> p((size_t) coo.n_cols + 1), i((size_t) (coo.nnz > 0 ? coo.nnz : 1));
    detail_cv::regroup<T>('C', coo, coo.n_cols, p.data(), i.data(), csc.vals, false);
    detail_cv::narrow(p, csc.colptr); i.resize((size_t) coo.nnz); detail_cv::narrow(i, csc.rowidxs);
}
template <typename T, typename sint_t1 = int64_t, typename sint_t2 = int64_t>
inline void coo_to_csr(COOMatrix<T, sint_t1>& coo, CSRMatrix<T, sint_t2>& csr) {
    randblas_require(csr.n_rows == coo.n_rows); randblas_require(csr.n_cols == coo.n_cols);
    randblas_require(csr.index_base == IndexBase::Zero); randblas_require(coo.index_base == IndexBase::Zero);
    csr.reserve(coo.nnz);
    if (std::is_same<sint_t2, int64_t>::value) { detail_cv::regroup<T>('R', coo, coo.n_rows, (int64_t*) csr.rowptr, (int64_t*) csr.colidxs, csr.vals, csr.on_device); return; }
    randblas_require(!csr.on_device);
    std::vector<int64_t> p((size_t) coo.n_rows + 1), i((size_t) (coo.nnz > 0 ? coo.nnz : 1));
    detail_cv::regroup<T>('R', coo, coo.n_rows, p.data(), i.data(), csr.vals, false);
    detail_cv::narrow(p, csr.rowptr); i.resize((size_t) coo.nnz); detail_cv::narrow(i, csr.colidxs);
}
// CSC / CSR -> COO (conversions.hh:64-82, 105-123). The reference writes loop counters where the stored indices belong
// (:75, :117); the evident intent -- row / column index of every stored entry -- is what is produced here. Host arrays only.
template <typename T, typename sint_t1 = int64_t, typename sint_t2 = int64_t>
inline void csc_to_coo(CSCMatrix<T, sint_t1>& csc, COOMatrix<T, sint_t2>& coo) {
    randblas_require(csc.n_rows == coo.n_rows); randblas_require(csc.n_cols == coo.n_cols);
    randblas_require(csc.index_base == IndexBase::Zero); randblas_require(coo.index_base == IndexBase::Zero);
    randblas_require(!device_pointers());
    coo.reserve(csc.nnz);
    for (int64_t j = 0; j < csc.n_cols; ++j)
        for (sint_t1 e = csc.colptr[j]; e < csc.colptr[j + 1]; ++e) { coo.vals[e] = csc.vals[e]; coo.rows[e] = (sint_t2) csc.rowidxs[e]; coo.cols[e] = (sint_t2) j; }
    coo.sort = NonzeroSort::CSC;
}
template <typename T, typename sint_t1 = int64_t, typename sint_t2 = int64_t>
inline void csr_to_coo(CSRMatrix<T, sint_t1>& csr, COOMatrix<T, sint_t2>& coo) {
    randblas_require(csr.n_rows == coo.n_rows); randblas_require(csr.n_cols == coo.n_cols);
    randblas_require(csr.index_base == IndexBase::Zero); randblas_require(coo.index_base == IndexBase::Zero);
    randblas_require(!device_pointers());
    coo.reserve(csr.nnz);
    for (int64_t i = 0; i < csr.n_rows; ++i)
        for (sint_t1 e = csr.rowptr[i]; e < csr.rowptr[i + 1]; ++e) { coo.vals[e] = csr.vals[e]; coo.rows[e] = (sint_t2) i; coo.cols[e] = (sint_t2) csr.colidxs[e]; }
    coo.sort = NonzeroSort::CSR;
}
// zero-copy transposes (conversions.hh:125-159); share_memory = false copies on the host
template <typename T>
inline CSRMatrix<T> transpose_as_csr(CSCMatrix<T>& A, bool share_memory = true) {
    if (share_memory) return CSRMatrix<T>(A.n_cols, A.n_rows, A.nnz, A.vals, A.colptr, A.rowidxs);
    randblas_require(!device_pointers());
    CSRMatrix<T> At(A.n_cols, A.n_rows);
    At.reserve(A.nnz);
    std::copy(A.rowidxs, A.rowidxs + A.nnz, At.colidxs); std::copy(A.vals, A.vals + A.nnz, At.vals); std::copy(A.colptr, A.colptr + A.n_cols + 1, At.rowptr);
    return At;
}
template <typename T>
inline CSCMatrix<T> transpose_as_csc(CSRMatrix<T>& A, bool share_memory = true) {
    if (share_memory) return CSCMatrix<T>(A.n_cols, A.n_rows, A.nnz, A.vals, A.colidxs, A.rowptr);
    randblas_require(!device_pointers());
    CSCMatrix<T> At(A.n_cols, A.n_rows);
    At.reserve(A.nnz);
    std::copy(A.colidxs, A.colidxs + A.nnz, At.rowidxs); std::copy(A.vals, A.vals + A.nnz, At.vals); std::copy(A.rowptr, A.rowptr + A.n_rows + 1, At.colptr);
    return At;
}
}  // namespace conversions
}  // namespace sparse_data
using sparse_data::left_spmm; using sparse_data::right_spmm; using sparse_data::spmm;

}  // namespace RandBLAS
