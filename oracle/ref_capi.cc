// ORACLE / TEST INFRASTRUCTURE ONLY -- never part of the shipped product path.
//
// extern "C" wrappers that instantiate the UNMODIFIED reference templates from
// /root/reference/RandBLAS.hh (compiled where they lie, against oracle/shims/) so that
// tests and bench.py's cpu_baseline leg can call the real reference through ctypes.
// Built by oracle/Makefile into oracle/_ref/librandblas_ref.so (git-ignored; travels
// to the GPU box with the snapshot). No reference source is copied into this repo.
//
// Conventions: layout 'C'/'R'; op 'N'/'T'; family 'G'/'U'; major axis 'S'/'L';
// state = uint32 ctr[4] + uint32 key[2]. Every function returns 0 on success,
// 1 if the reference threw RandBLAS::exceptions::Error, 2 for any other exception.
#include <RandBLAS.hh>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>

using namespace RandBLAS;
using RNG = r123::Philox4x32;
using State = RNGState<RNG>;

static thread_local std::string g_last_error;

static State mk_state(const uint32_t* ctr, const uint32_t* key) {
    State s;
    for (int i = 0; i < 4; ++i) s.counter.v[i] = ctr[i];
    for (int i = 0; i < 2; ++i) s.key.v[i] = key[i];
    return s;
}
static void put_ctr(const State& s, uint32_t* out) {
    if (!out) return;
    for (int i = 0; i < 4; ++i) out[i] = s.counter.v[i];
}
static blas::Layout L(char c) { return c == 'C' ? blas::Layout::ColMajor : blas::Layout::RowMajor; }
static blas::Op O(char c) { return c == 'N' ? blas::Op::NoTrans : blas::Op::Trans; }
static DenseDist mk_ddist(int64_t r, int64_t c, char fam, char ma) {
    return DenseDist(r, c, (DenseDistName) fam, (MajorAxis) ma);
}

#define REF_TRY try {
#define REF_CATCH                                                                         \
    } catch (const RandBLAS::exceptions::Error& e) { g_last_error = e.what(); return 1; } \
      catch (const std::exception& e) { g_last_error = e.what(); return 2; }              \
    return 0;

extern "C" {

const char* ref_last_error() { return g_last_error.c_str(); }
const char* ref_blas_backend() { return blas_shim::backend_name(); }
int ref_blas_threads() {
    auto& be = blas_shim::backend();
    return be.get_threads ? be.get_threads() : omp_get_max_threads();
}
int ref_omp_threads() { return omp_get_max_threads(); }
// bench.py: torch.distributed.run exports OMP_NUM_THREADS=1; the CPU legs ask for every host core explicitly
int ref_set_threads(int n) {
    if (n < 1) n = 1;
    omp_set_num_threads(n);
    auto& be = blas_shim::backend();
    if (be.set_threads) be.set_threads(n);
    return omp_get_max_threads();
}

// ---- raw generator pieces (Random123 shim + random_gen.hh) --------------------------
int ref_philox4x32(const uint32_t* ctr, const uint32_t* key, uint32_t* out) {
    REF_TRY
    State s = mk_state(ctr, key);
    RNG g;
    auto r = g(s.counter, s.key);
    for (int i = 0; i < 4; ++i) out[i] = r.v[i];
    REF_CATCH
}
int ref_ctr_incr(uint32_t* ctr, uint64_t n) {
    REF_TRY
    RNG::ctr_type c;
    for (int i = 0; i < 4; ++i) c.v[i] = ctr[i];
    c.incr(n);
    for (int i = 0; i < 4; ++i) ctr[i] = c.v[i];
    REF_CATCH
}
int ref_generate4(char family, const uint32_t* ctr, const uint32_t* key, float* out) {
    REF_TRY
    State s = mk_state(ctr, key);
    RNG g;
    if (family == 'G') { auto r = r123ext::boxmul::generate(g, s.counter, s.key); for (int i = 0; i < 4; ++i) out[i] = r[i]; }
    else               { auto r = r123ext::uneg11::generate(g, s.counter, s.key); for (int i = 0; i < 4; ++i) out[i] = r[i]; }
    REF_CATCH
}
// transform applied to caller-chosen raw words (for exhaustive libm sweeps)
int ref_boxmuller_raw(int64_t n, const uint32_t* u0, const uint32_t* u1, float* s_out, float* c_out) {
    REF_TRY
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        auto f = r123::boxmuller(u0[i], u1[i]);
        s_out[i] = f.x; c_out[i] = f.y;
    }
    REF_CATCH
}

// ---- dense operators (dense_skops.hh) -----------------------------------------------
#define DEF_FILL_SUBMAT(SUF, T)                                                                          \
int ref_fill_dense_submat_##SUF(char family, int64_t n_cols, T* smat, int64_t n_srows, int64_t n_scols,   \
                                 int64_t ptr, const uint32_t* ctr, const uint32_t* key, int64_t lda,        \
                                 uint32_t* next_ctr) {                                                      \
    REF_TRY                                                                                                 \
    State s = mk_state(ctr, key);                                                                           \
    State nx = (family == 'G')                                                                              \
        ? dense::fill_dense_submat_impl<T, RNG, r123ext::boxmul>(n_cols, smat, n_srows, n_scols, ptr, s, lda) \
        : dense::fill_dense_submat_impl<T, RNG, r123ext::uneg11>(n_cols, smat, n_srows, n_scols, ptr, s, lda); \
    put_ctr(nx, next_ctr);                                                                                  \
    REF_CATCH                                                                                               \
}
DEF_FILL_SUBMAT(f32, float)
DEF_FILL_SUBMAT(f64, double)

#define DEF_FILL_DENSE(SUF, T)                                                                            \
int ref_fill_dense_##SUF(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,        \
                          int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, T* buff,              \
                          const uint32_t* ctr, const uint32_t* key, uint32_t* next_ctr) {                   \
    REF_TRY                                                                                                 \
    DenseDist D = mk_ddist(D_rows, D_cols, family, major_axis);                                             \
    State nx = fill_dense(L(layout), D, n_rows, n_cols, ro_s, co_s, buff, mk_state(ctr, key));              \
    put_ctr(nx, next_ctr);                                                                                  \
    REF_CATCH                                                                                               \
}
DEF_FILL_DENSE(f32, float)
DEF_FILL_DENSE(f64, double)

int ref_dense_dist_info(int64_t D_rows, int64_t D_cols, char family, char major_axis,
                        const uint32_t* ctr, const uint32_t* key,
                        char* layout_out, int64_t* ma_len_out, uint32_t* next_ctr) {
    REF_TRY
    DenseDist D = mk_ddist(D_rows, D_cols, family, major_axis);
    *layout_out = (char) dist_to_layout(D);
    *ma_len_out = major_axis_length(D);
    DenseSkOp<float> S(D, mk_state(ctr, key));
    put_ctr(S.next_state, next_ctr);
    REF_CATCH
}

// ---- sparse operators (sparse_skops.hh) ---------------------------------------------
#define DEF_FILL_SPARSE(SUF, T, I)                                                                        \
int ref_fill_sparse_##SUF(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis,                \
                           const uint32_t* ctr, const uint32_t* key, I* rows, I* cols, T* vals,             \
                           uint32_t* next_ctr) {                                                            \
    REF_TRY                                                                                                 \
    SparseDist D = {.n_rows = n_rows, .n_cols = n_cols, .vec_nnz = vec_nnz, .major_axis = (MajorAxis) major_axis}; \
    SparseSkOp<T, RNG, I> S(D, mk_state(ctr, key), rows, cols, vals, false);                                \
    fill_sparse(S);                                                                                         \
    put_ctr(S.next_state, next_ctr);                                                                        \
    REF_CATCH                                                                                               \
}
DEF_FILL_SPARSE(f32_i32, float, int)
DEF_FILL_SPARSE(f32_i64, float, int64_t)
DEF_FILL_SPARSE(f64_i32, double, int)
DEF_FILL_SPARSE(f64_i64, double, int64_t)

int ref_repeated_fisher_yates_i64(const uint32_t* ctr, const uint32_t* key, int64_t k, int64_t n, int64_t r,
                                  int64_t* idxs) {
    REF_TRY
    sparse::repeated_fisher_yates(mk_state(ctr, key), k, n, r, idxs);
    REF_CATCH
}

// ---- sketch_general with a dense operator: lskge3 / rskge3 (skge.hh:173-364) ----------
// side 'L': B(d x n) = alpha*op(submat(S))*op(A) + beta*B ; dims (d, n, m)
// side 'R': B(m x d) = alpha*op(A)*op(submat(S)) + beta*B ; dims passed as (m, d, n) in (p0,p1,p2)
// If S_buff != nullptr it is attached as S.buff with layout S_layout (pre-filled operator).
#define DEF_SK_DENSE(SUF, T)                                                                              \
int ref_sketch_dense_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2,  \
                            T alpha, int64_t D_rows, int64_t D_cols, char family, char major_axis,           \
                            const uint32_t* ctr, const uint32_t* key, T* S_buff, char S_layout,              \
                            int64_t ro_s, int64_t co_s, const T* A, int64_t lda, T beta, T* B, int64_t ldb) { \
    REF_TRY                                                                                                 \
    DenseDist D = mk_ddist(D_rows, D_cols, family, major_axis);                                             \
    DenseSkOp<T> S(D, mk_state(ctr, key));                                                                  \
    if (S_buff) { S.buff = S_buff; S.layout = L(S_layout); }                                                \
    if (side == 'L') sketch_general(L(layout), O(opS), O(opA), p0, p1, p2, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); \
    else             sketch_general(L(layout), O(opA), O(opS), p0, p1, p2, alpha, A, lda, S, ro_s, co_s, beta, B, ldb); \
    REF_CATCH                                                                                               \
}
DEF_SK_DENSE(f32, float)
DEF_SK_DENSE(f64, double)

// ---- sketch_general with a sparse operator: lskges / rskges (skge.hh:485-641) ---------
#define DEF_SK_SPARSEOP(SUF, T)                                                                           \
int ref_sketch_sparseop_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, \
                               T alpha, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis,     \
                               const uint32_t* ctr, const uint32_t* key, int64_t ro_s, int64_t co_s,          \
                               const T* A, int64_t lda, T beta, T* B, int64_t ldb, double* t_fill, double* t_apply) { \
    REF_TRY                                                                                                 \
    SparseDist D = {.n_rows = S_rows, .n_cols = S_cols, .vec_nnz = vec_nnz, .major_axis = (MajorAxis) major_axis}; \
    SparseSkOp<T> S(D, mk_state(ctr, key));                                                                 \
    auto t0 = std::chrono::steady_clock::now();                                                             \
    fill_sparse(S);                                                                                         \
    auto t1 = std::chrono::steady_clock::now();                                                             \
    if (side == 'L') sketch_general(L(layout), O(opS), O(opA), p0, p1, p2, alpha, S, ro_s, co_s, A, lda, beta, B, ldb); \
    else             sketch_general(L(layout), O(opA), O(opS), p0, p1, p2, alpha, A, lda, S, ro_s, co_s, beta, B, ldb); \
    auto t2 = std::chrono::steady_clock::now();                                                             \
    if (t_fill)  *t_fill  = std::chrono::duration<double>(t1 - t0).count();                                 \
    if (t_apply) *t_apply = std::chrono::duration<double>(t2 - t1).count();                                 \
    REF_CATCH                                                                                               \
}
DEF_SK_SPARSEOP(f32, float)
DEF_SK_SPARSEOP(f64, double)

// ---- sketch_sparse: dense operator x sparse data, lsksp3 / rsksp3 (sksp.hh:147-345) ---
// fmt 'O' = COO (idx0 = rows, idx1 = cols), 'R' = CSR (idx0 = rowptr, idx1 = colidxs),
// 'C' = CSC (idx0 = rowidxs, idx1 = colptr). Indices int64.
}  // extern "C" (templates need C++ linkage)
template <typename T, typename SpMat>
static void sk_sparse_dispatch(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2,
                               T alpha, DenseSkOp<T>& S, int64_t ro_s, int64_t co_s, SpMat& A,
                               int64_t ro_a, int64_t co_a, T beta, T* B, int64_t ldb) {
    if (side == 'L') sketch_sparse(L(layout), O(opS), O(opA), p0, p1, p2, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb);
    else             sketch_sparse(L(layout), O(opA), O(opS), p0, p1, p2, alpha, A, ro_a, co_a, S, ro_s, co_s, beta, B, ldb);
}
extern "C" {
#define DEF_SK_SPARSEDATA(SUF, T)                                                                         \
int ref_sketch_sparsedata_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, \
                                 T alpha, int64_t D_rows, int64_t D_cols, char family, char major_axis,       \
                                 const uint32_t* ctr, const uint32_t* key, int64_t ro_s, int64_t co_s,        \
                                 char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz, T* vals,              \
                                 int64_t* idx0, int64_t* idx1, int64_t ro_a, int64_t co_a,                    \
                                 T beta, T* B, int64_t ldb) {                                                 \
    REF_TRY                                                                                                 \
    DenseDist D = mk_ddist(D_rows, D_cols, family, major_axis);                                             \
    DenseSkOp<T> S(D, mk_state(ctr, key));                                                                  \
    if (fmt == 'O') { sparse_data::COOMatrix<T> A(A_rows, A_cols, nnz, vals, idx0, idx1);                   \
        sk_sparse_dispatch(side, layout, opS, opA, p0, p1, p2, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb); } \
    else if (fmt == 'R') { sparse_data::CSRMatrix<T> A(A_rows, A_cols, nnz, vals, idx0, idx1);              \
        sk_sparse_dispatch(side, layout, opS, opA, p0, p1, p2, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb); } \
    else { sparse_data::CSCMatrix<T> A(A_rows, A_cols, nnz, vals, idx0, idx1);                              \
        sk_sparse_dispatch(side, layout, opS, opA, p0, p1, p2, alpha, S, ro_s, co_s, A, ro_a, co_a, beta, B, ldb); } \
    REF_CATCH                                                                                               \
}
DEF_SK_SPARSEDATA(f32, float)
DEF_SK_SPARSEDATA(f64, double)

// ---- wrappers: sketch_vector (skve.hh:152-176), sketch_symmetric (sksy.hh:165-319) ----
#define DEF_SK_VECTOR(SUF, T)                                                                             \
int ref_sketch_vector_dense_##SUF(char opS, int64_t d, int64_t m, T alpha, int64_t D_rows, int64_t D_cols,  \
                                   char family, char major_axis, const uint32_t* ctr, const uint32_t* key,   \
                                   int64_t ro_s, int64_t co_s, const T* x, int64_t incx, T beta, T* y, int64_t incy) { \
    REF_TRY                                                                                                 \
    DenseSkOp<T> S(mk_ddist(D_rows, D_cols, family, major_axis), mk_state(ctr, key));                       \
    sketch_vector(O(opS), d, m, alpha, S, ro_s, co_s, x, incx, beta, y, incy);                              \
    REF_CATCH                                                                                               \
}
DEF_SK_VECTOR(f32, float)
DEF_SK_VECTOR(f64, double)

#define DEF_SK_SYMM(SUF, T)                                                                               \
int ref_sketch_symmetric_dense_##SUF(char side, char layout, int64_t p0, int64_t p1, T alpha,               \
                                      int64_t D_rows, int64_t D_cols, char family, char major_axis,          \
                                      const uint32_t* ctr, const uint32_t* key, int64_t ro_s, int64_t co_s,  \
                                      const T* A, int64_t lda, T beta, T* B, int64_t ldb, T sym_check_tol) { \
    REF_TRY                                                                                                 \
    DenseSkOp<T> S(mk_ddist(D_rows, D_cols, family, major_axis), mk_state(ctr, key));                       \
    if (side == 'L') sketch_symmetric(L(layout), p0, p1, alpha, S, ro_s, co_s, A, lda, beta, B, ldb, sym_check_tol); \
    else             sketch_symmetric(L(layout), p0, p1, alpha, A, lda, S, ro_s, co_s, beta, B, ldb, sym_check_tol); \
    REF_CATCH                                                                                               \
}
DEF_SK_SYMM(f32, float)
DEF_SK_SYMM(f64, double)

}  // extern "C"

// ---- index sampling utilities (util.hh:257-334) ----------------------------------------
extern "C" {

int ref_sample_indices_iid_uniform(int64_t n, int64_t k, int64_t* samples, const uint32_t* ctr, const uint32_t* key,
                                   uint32_t* next_ctr) {
    REF_TRY
    State out = RandBLAS::util::sample_indices_iid_uniform(n, k, samples, mk_state(ctr, key));
    put_ctr(out, next_ctr);
    REF_CATCH
}
#define DEF_SAMPLE(SUF, T)                                                                                  \
int ref_sample_indices_iid_##SUF(int64_t n, T* cdf, int64_t k, int64_t* samples, const uint32_t* ctr,        \
                                 const uint32_t* key, uint32_t* next_ctr) {                                  \
    REF_TRY                                                                                                 \
    State out = RandBLAS::util::sample_indices_iid(n, cdf, k, samples, mk_state(ctr, key));                  \
    put_ctr(out, next_ctr);                                                                                 \
    REF_CATCH                                                                                               \
}                                                                                                           \
int ref_weights_to_cdf_##SUF(int64_t n, T* w, T error_if_below) {                                           \
    REF_TRY                                                                                                 \
    RandBLAS::util::weights_to_cdf(n, w, error_if_below);                                                   \
    REF_CATCH                                                                                               \
}
DEF_SAMPLE(f32, float)
DEF_SAMPLE(f64, double)

}  // extern "C"
