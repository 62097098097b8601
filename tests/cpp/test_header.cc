// Exercises include/RandBLAS_b200.hh the way a RandBLAS caller would (host pointers, the reference's argument order),
// and checks every result against the CPU oracle (oracle/randblas_oracle.c -- test infrastructure, linked here only as
// the checker). Built with g++ (no nvcc, no CUDA headers) by tests/test_cpp_header.py; runs on a GPU box.
//
// Mirrors the structure of the reference's own tests: test/test_datastructures/test_denseskop.cc (bit-exact fills,
// next_state), test_sparseskop.cc (fill_sparse), test/test_matmul_cores/test_lskge3.cc, test_rskge3.cc, test_lskges.cc
// (sketches against an explicit reference product within the comparison.hh bound).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>
#include <sstream>
#include <RandBLAS.hh>                       // the drop-in name: include/RandBLAS.hh forwards to RandBLAS_b200.hh

extern "C" {
int orc_fill_dense_f32(char, int64_t, int64_t, char, char, int64_t, int64_t, int64_t, int64_t, float*, const uint32_t*, const uint32_t*, uint32_t*);
int orc_fill_dense_f64(char, int64_t, int64_t, char, char, int64_t, int64_t, int64_t, int64_t, double*, const uint32_t*, const uint32_t*, uint32_t*);
int orc_fill_sparse_f64_i64(int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*, int64_t*, int64_t*, double*, uint32_t*);
int orc_fill_sparse_f32_i32(int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*, int32_t*, int32_t*, float*, uint32_t*);
int orc_sketch_dense_f32(char, char, char, char, int64_t, int64_t, int64_t, float, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                         int64_t, int64_t, const float*, int64_t, float, float*, int64_t);
int orc_sketch_dense_f64(char, char, char, char, int64_t, int64_t, int64_t, double, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                         int64_t, int64_t, const double*, int64_t, double, double*, int64_t);
int orc_sketch_sparseop_f64(char, char, char, char, int64_t, int64_t, int64_t, double, int64_t, int64_t, int64_t, char, const uint32_t*, const uint32_t*,
                            int64_t, int64_t, const double*, int64_t, double, double*, int64_t);
int orc_sketch_sparsedata_f64(char, char, char, char, int64_t, int64_t, int64_t, double, int64_t, int64_t, char, char, const uint32_t*, const uint32_t*,
                              int64_t, int64_t, char, int64_t, int64_t, int64_t, const double*, const int64_t*, const int64_t*, int64_t, int64_t,
                              double, double*, int64_t);
int orc_sample_indices_iid_uniform(int64_t, int64_t, int64_t*, const uint32_t*, const uint32_t*, uint32_t*);
int orc_sample_indices_iid_f64(int64_t, const double*, int64_t, int64_t*, const uint32_t*, const uint32_t*, uint32_t*);
int orc_weights_to_cdf_f64(int64_t, double*, double);
}

using namespace RandBLAS;
static int failures = 0;
#define CHECK(cond) do { if (!(cond)) { std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); ++failures; } } while (0)

template <typename T> static bool bits_equal(const std::vector<T>& a, const std::vector<T>& b) {
    return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(T)) == 0;
}
template <typename T> static double max_rel(const std::vector<T>& a, const std::vector<T>& b) {
    double num = 0, den = 1e-300;
    for (size_t i = 0; i < a.size(); ++i) { num = std::max(num, (double) std::abs(a[i] - b[i])); den = std::max(den, (double) std::abs(b[i])); }
    return num / den;
}

static void test_fill_dense() {
    for (auto fam : { DenseDistName::Gaussian, DenseDistName::Uniform })
        for (auto ma : { MajorAxis::Long, MajorAxis::Short })
            for (int64_t cols : { 10, 11, 12, 13 }) {            // all four n_cols % 4 phases
                DenseDist D(7, cols, fam, ma);
                RNGState<> seed(42);
                std::vector<float> got(7 * cols), want(7 * cols);
                RNGState<> next = fill_dense(D, got.data(), seed);
                uint32_t nx[4];
                CHECK(orc_fill_dense_f32(detail::ch(dist_to_layout(D)), 7, cols, (char) fam, (char) ma, 7, cols, 0, 0, want.data(), seed.counter.v, seed.key.v, nx) == 0);
                CHECK(bits_equal(got, want));
                CHECK(std::memcmp(next.counter.v, nx, 16) == 0);
                DenseSkOp<float> S(D, seed);
                CHECK(std::memcmp(S.next_state.counter.v, nx, 16) == 0);
                // sub-matrix with offsets, both layouts, double
                std::vector<double> g2(3 * 5), w2(3 * 5);
                for (auto lay : { blas::Layout::ColMajor, blas::Layout::RowMajor }) {
                    fill_dense(lay, D, 3, 5, 2, 4, g2.data(), seed);
                    CHECK(orc_fill_dense_f64(detail::ch(lay), 7, cols, (char) fam, (char) ma, 3, 5, 2, 4, w2.data(), seed.counter.v, seed.key.v, nx) == 0);
                    CHECK(bits_equal(g2, w2));
                }
            }
    // fill_dense(DenseSkOp&) allocates, fills and owns the buffer
    DenseSkOp<double> S(DenseDist(6, 40), 7);
    fill_dense(S);
    CHECK(S.buff != nullptr && S.del_buff_on_destruct);
    std::vector<double> want(240), got(S.buff, S.buff + 240);
    uint32_t nx[4];
    orc_fill_dense_f64(detail::ch(S.layout), 6, 40, 'G', 'L', 6, 40, 0, 0, want.data(), S.seed_state.counter.v, S.seed_state.key.v, nx);
    CHECK(bits_equal(got, want));
    bool threw = false;
    try { std::vector<float> b(4); fill_dense(DenseDist(2, 2, DenseDistName::BlackBox), b.data(), RNGState<>(0)); } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw);
}

static void test_fill_sparse() {
    for (auto ma : { MajorAxis::Short, MajorAxis::Long })
        for (auto shape : { std::pair<int64_t, int64_t>{ 7, 20 }, { 20, 7 }, { 15, 15 } }) {
            SparseDist D{ shape.first, shape.second, 3, ma };
            SparseSkOp<double> S(D, 99);
            fill_sparse(S);
            const int64_t n = nnz(S);
            std::vector<int64_t> r(n), c(n); std::vector<double> v(n); uint32_t nx[4];
            CHECK(orc_fill_sparse_f64_i64(D.n_rows, D.n_cols, 3, (char) ma, S.seed_state.counter.v, S.seed_state.key.v, r.data(), c.data(), v.data(), nx) == 0);
            CHECK(std::memcmp(r.data(), S.rows, n * 8) == 0 && std::memcmp(c.data(), S.cols, n * 8) == 0 && std::memcmp(v.data(), S.vals, n * 8) == 0);
            CHECK(std::memcmp(nx, S.next_state.counter.v, 16) == 0);
            SparseSkOp<float, r123::Philox4x32, int32_t> S32(D, 99);
            fill_sparse(S32);
            std::vector<int32_t> r32(n), c32(n); std::vector<float> v32(n);
            CHECK(orc_fill_sparse_f32_i32(D.n_rows, D.n_cols, 3, (char) ma, S32.seed_state.counter.v, S32.seed_state.key.v, r32.data(), c32.data(), v32.data(), nx) == 0);
            CHECK(std::memcmp(r32.data(), S32.rows, n * 4) == 0 && std::memcmp(c32.data(), S32.cols, n * 4) == 0 && std::memcmp(v32.data(), S32.vals, n * 4) == 0);
        }
}

static std::vector<double> randmat(int64_t n, uint32_t key) {
    std::vector<double> a(n);
    fill_dense(DenseDist(n, 1), a.data(), RNGState<>(key));
    return a;
}

static void test_sketch_dense() {
    const int64_t d = 9, n = 13, m = 50;
    DenseSkOp<double> S(DenseDist(d, m), 0);
    std::vector<double> A = randmat(m * n, 99);
    for (auto lay : { blas::Layout::ColMajor, blas::Layout::RowMajor }) {
        const int64_t lda = lay == blas::Layout::ColMajor ? m : n, ldb = lay == blas::Layout::ColMajor ? d : n;
        std::vector<double> B(d * n, 0.5), W(d * n, 0.5);
        sketch_general(lay, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.5, S, A.data(), lda, -0.5, B.data(), ldb);
        CHECK(orc_sketch_dense_f64('L', detail::ch(lay), 'N', 'N', d, n, m, 1.5, d, m, 'G', 'L', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, A.data(), lda, -0.5, W.data(), ldb) == 0);
        CHECK(max_rel(B, W) < 1e-13);
        // right sketch with the transposed operator: B (n x d) = A^T (n x m) * S^T (m x d)
        const int64_t ldb2 = lay == blas::Layout::ColMajor ? n : d;
        std::vector<double> B2(n * d, 0.0), W2(n * d, 0.0);
        sketch_general(lay, blas::Op::Trans, blas::Op::Trans, n, d, m, 1.0, A.data(), lda, S, 0.0, B2.data(), ldb2);
        CHECK(orc_sketch_dense_f64('R', detail::ch(lay), 'T', 'T', n, d, m, 1.0, d, m, 'G', 'L', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, A.data(), lda, 0.0, W2.data(), ldb2) == 0);
        CHECK(max_rel(B2, W2) < 1e-13);
    }
    // explicit buffer == implicit operator
    DenseSkOp<double> S2(DenseDist(d, m), 0);
    fill_dense(S2);
    std::vector<double> B(d * n, 0.0), W(d * n, 0.0);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S2, A.data(), m, 0.0, B.data(), d);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, A.data(), m, 0.0, W.data(), d);
    CHECK(max_rel(B, W) < 1e-13);
    // sketch_vector and dimension checks
    std::vector<double> x = randmat(m, 5), y(d, 0.0), wy(d, 0.0);
    sketch_vector(blas::Op::NoTrans, 1.0, S, x.data(), 1, 0.0, y.data(), 1);
    orc_sketch_dense_f64('L', 'R', 'N', 'N', d, 1, m, 1.0, d, m, 'G', 'L', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, x.data(), 1, 0.0, wy.data(), 1);
    CHECK(max_rel(y, wy) < 1e-13);
    bool threw = false;
    try { sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d + 1, n, m, 1.0, S, A.data(), m, 0.0, B.data(), d + 1); } catch (const Error&) { threw = true; }
    CHECK(threw);
    threw = false;
    try { sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, 0, 0, A.data(), m - 1, 0.0, B.data(), d); } catch (const Error&) { threw = true; }
    CHECK(threw);
    // float
    DenseSkOp<float> Sf(DenseDist(d, m, DenseDistName::Uniform), 3);
    std::vector<float> Af(m * n), Bf(d * n, 0.f), Wf(d * n, 0.f);
    fill_dense(DenseDist(m, n), Af.data(), RNGState<>(99));
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.f, Sf, Af.data(), m, 0.f, Bf.data(), d);
    orc_sketch_dense_f32('L', 'C', 'N', 'N', d, n, m, 1.f, d, m, 'U', 'L', Sf.seed_state.counter.v, Sf.seed_state.key.v, 0, 0, Af.data(), m, 0.f, Wf.data(), d);
    CHECK(max_rel(Bf, Wf) < 1e-5);
}

static void test_sketch_sparseop() {
    const int64_t d = 64, n = 24, m = 700;
    std::vector<double> A = randmat(m * n, 99);
    for (auto lay : { blas::Layout::ColMajor, blas::Layout::RowMajor }) {
        const int64_t lda = lay == blas::Layout::ColMajor ? m : n, ldb = lay == blas::Layout::ColMajor ? d : n;
        SparseSkOp<double> S({ d, m, 8, MajorAxis::Short }, 1997);
        std::vector<double> B(d * n, 0.0), W(d * n, 0.0), B2(d * n, 0.0);
        sketch_general(lay, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, A.data(), lda, 0.0, B.data(), ldb);     // implicit (streaming kernel)
        CHECK(orc_sketch_sparseop_f64('L', detail::ch(lay), 'N', 'N', d, n, m, 1.0, d, m, 8, 'S', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, A.data(), lda, 0.0, W.data(), ldb) == 0);
        CHECK(max_rel(B, W) < 1e-13);
        fill_sparse(S);                                                                                                       // explicit triplets
        sketch_general(lay, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, A.data(), lda, 0.0, B2.data(), ldb);
        CHECK(max_rel(B2, W) < 1e-13);
    }
    // symmetric wrapper: A symmetric n x n, right sketch by an n x d LASO
    const int64_t ns = 40, ds = 6;
    std::vector<double> G = randmat(ns * ns, 3), Sym(ns * ns);
    for (int64_t i = 0; i < ns; ++i) for (int64_t j = 0; j < ns; ++j) Sym[i + j * ns] = G[i + j * ns] + G[j + i * ns];
    SparseSkOp<double> SL({ ns, ds, 2, MajorAxis::Long }, 5);
    std::vector<double> B(ns * ds, 0.0), W(ns * ds, 0.0);
    sketch_symmetric(blas::Layout::ColMajor, 1.0, Sym.data(), ns, SL, 0.0, B.data(), ns);
    orc_sketch_sparseop_f64('R', 'C', 'N', 'N', ns, ds, ns, 1.0, ns, ds, 2, 'L', SL.seed_state.counter.v, SL.seed_state.key.v, 0, 0, Sym.data(), ns, 0.0, W.data(), ns);
    CHECK(max_rel(B, W) < 1e-13);
    Sym[1] += 1.0;
    bool threw = false;
    try { sketch_symmetric(blas::Layout::ColMajor, 1.0, Sym.data(), ns, SL, 0.0, B.data(), ns); } catch (const Error&) { threw = true; }
    CHECK(threw);
}

static void test_sketch_sparse() {
    // CSR data 30 x 20 with a few nonzeros per row, left sketch with a 5 x 30 Gaussian operator
    const int64_t m = 30, n = 20, d = 5;
    CSRMatrix<double> A(m, n);
    A.reserve(3 * m);
    std::vector<double> vals = randmat(3 * m, 8);
    for (int64_t i = 0; i <= m; ++i) A.rowptr[i] = 3 * i;
    for (int64_t i = 0; i < m; ++i) for (int k = 0; k < 3; ++k) { A.colidxs[3 * i + k] = (i * 7 + k * 5) % n; A.vals[3 * i + k] = vals[3 * i + k]; }
    DenseSkOp<double> S(DenseDist(d, m), 0);
    std::vector<double> B(d * n, 0.0), W(d * n, 0.0);
    sketch_sparse(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, 0, 0, A, 0, 0, 0.0, B.data(), d);
    CHECK(orc_sketch_sparsedata_f64('L', 'C', 'N', 'N', d, n, m, 1.0, d, m, 'G', 'L', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, 'R', m, n, A.nnz,
                                    A.vals, A.rowptr, A.colidxs, 0, 0, 0.0, W.data(), d) == 0);
    CHECK(max_rel(B, W) < 1e-13);
}

static void test_index_sampling() {
    // util::weights_to_cdf / sample_indices_iid / sample_indices_iid_uniform (util.hh:257-334), host pointers
    const int64_t n = 500, k = 1001;
    std::vector<double> w = randmat(n, 4), cdf(n);
    for (auto& v : w) v = std::abs(v);
    cdf = w;
    util::weights_to_cdf(n, w.data());
    CHECK(orc_weights_to_cdf_f64(n, cdf.data(), -std::numeric_limits<double>::epsilon()) == 0);
    CHECK(bits_equal(w, cdf));
    RNGState<> st(11);
    std::vector<int64_t> got(k), want(k);
    uint32_t next[4];
    RNGState<> out = util::sample_indices_iid(n, w.data(), k, got.data(), st);
    orc_sample_indices_iid_f64(n, cdf.data(), k, want.data(), st.counter.v, st.key.v, next);
    CHECK(got == want && std::memcmp(out.counter.v, next, 16) == 0);
    out = util::sample_indices_iid_uniform(n, k, got.data(), st);
    orc_sample_indices_iid_uniform(n, k, want.data(), st.counter.v, st.key.v, next);
    CHECK(got == want && std::memcmp(out.counter.v, next, 16) == 0);
    bool threw = false;
    std::vector<double> bad = { 0.5, -0.1, 0.5 };
    try { util::weights_to_cdf(3, bad.data()); } catch (const Error&) { threw = true; }
    CHECK(threw);
}


// base.hh:161-251: RNGState with Random123-style arrays (what INTEGRATION.md's stub and reference callers touch)
static void test_rngstate_surface() {
    RNGState<> s(7);
    CHECK(s.key[0] == 7 && s.key.v[1] == 0 && s.counter.v[0] == 0 && RNGState<>::len_c == 4 && RNGState<>::len_k == 2);
    s.counter.v[0] = 0xFFFFFFFFu; s.counter.v[1] = 0xFFFFFFFFu;
    s.counter.incr(1);                                            // carry across two words (test_r123.cc:679-766)
    CHECK(s.counter[0] == 0 && s.counter[1] == 0 && s.counter[2] == 1 && s.counter[3] == 0);
    uint32_t c2[4] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0, 0 };
    rbx_ctr_add(c2, 1);
    CHECK(std::memcmp(c2, s.counter.v, 16) == 0);
    s.counter.incr(0x100000005ull);
    CHECK(s.counter[0] == 5 && s.counter[1] == 1 && s.counter[2] == 1);
    RNGState<> t = s;                                             // copies (base.hh:217-232)
    CHECK(t.counter == s.counter && t.key == s.key);
    s.key.incr(3);
    CHECK(s.key[0] == 10 && t.key[0] == 7);
    std::ostringstream os; os << t;
    CHECK(os.str() == "counter : {5, 1, 1, 0}\nkey     : {7, 0}");
}

// sparse_skops.hh:484-513 and spmm_dispatch.hh / conversions.hh: views, transposes, conversions and the public SpMM
static void test_sparse_surface() {
    const int64_t d = 12, m = 90, n = 7;
    SparseDist D = { d, m, 3, MajorAxis::Short };
    SparseSkOp<double> S(D, RNGState<>(5));
    auto Sv = coo_view_of_skop(S);                                // fills S
    CHECK(S.known_filled && Sv.nnz == 3 * m && Sv.rows == S.rows && !Sv.own_memory);
    auto St = transpose(S);
    CHECK(St.dist.n_rows == m && St.dist.n_cols == d && St.rows == S.cols && St.cols == S.rows && St.known_filled && !St.own_memory);
    std::vector<double> A(m * n), B1(d * n, 0.0), B2(d * n, 0.0), B3(d * n, 0.0), W(d * n, 0.0);
    for (size_t i = 0; i < A.size(); ++i) A[i] = std::sin(0.37 * i);
    orc_sketch_sparseop_f64('L', 'C', 'N', 'N', d, n, m, 1.0, d, m, 3, 'S', S.seed_state.counter.v, S.seed_state.key.v, 0, 0, A.data(), m, 0.0, W.data(), d);
    left_spmm(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, Sv, 0, 0, A.data(), m, 0.0, B1.data(), d);
    CHECK(max_rel(B1, W) < 1e-13);
    sketch_general(blas::Layout::ColMajor, blas::Op::Trans, blas::Op::NoTrans, d, n, m, 1.0, St, A.data(), m, 0.0, B2.data(), d);     // S^T transposed back
    CHECK(max_rel(B2, W) < 1e-13);
    // COO -> CSR / CSC on the device, then the same product through each format and through right_spmm
    sparse_data::CSRMatrix<double> R(d, m); sparse_data::CSCMatrix<double> C(d, m);
    sparse_data::conversions::coo_to_csr(Sv, R);
    sparse_data::conversions::coo_to_csc(Sv, C);
    CHECK(R.nnz == Sv.nnz && C.nnz == Sv.nnz && R.rowptr[0] == 0 && R.rowptr[d] == Sv.nnz && C.colptr[m] == Sv.nnz);
    bool sorted = true;
    for (int64_t i = 0; i < d; ++i) for (int64_t e = R.rowptr[i] + 1; e < R.rowptr[i + 1]; ++e) sorted = sorted && R.colidxs[e - 1] < R.colidxs[e];
    for (int64_t j = 0; j < m; ++j) { CHECK(C.colptr[j + 1] - C.colptr[j] == 3); for (int64_t e = C.colptr[j] + 1; e < C.colptr[j + 1]; ++e) sorted = sorted && C.rowidxs[e - 1] < C.rowidxs[e]; }
    CHECK(sorted);
    left_spmm(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, R, 0, 0, A.data(), m, 0.0, B3.data(), d);
    CHECK(max_rel(B3, W) < 1e-13);
    std::fill(B3.begin(), B3.end(), 0.0);
    spmm(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, C, 0, 0, A.data(), m, 0.0, B3.data(), d);
    CHECK(max_rel(B3, W) < 1e-13);
    auto Ct = sparse_data::conversions::transpose_as_csr(C);     // zero-copy: C^T as CSR (m x d)
    CHECK(Ct.n_rows == m && Ct.n_cols == d && Ct.rowptr == C.colptr && Ct.colidxs == C.rowidxs);
    // right_spmm: Bt (n x d) = At (n x m) * S^T (m x d), RowMajor buffers are the ColMajor ones transposed
    std::vector<double> Bt(n * d, 0.0);
    right_spmm(blas::Layout::RowMajor, blas::Op::NoTrans, blas::Op::NoTrans, n, d, m, 1.0, A.data(), m, Ct, 0, 0, 0.0, Bt.data(), d);
    CHECK(max_rel(Bt, W) < 1e-13);
    sparse_data::COOMatrix<double> back(d, m);
    sparse_data::conversions::csr_to_coo(R, back);
    CHECK(back.nnz == Sv.nnz && back.sort == sparse_data::NonzeroSort::CSR && back.rows[0] == 0);
    // host-pointer sketch into a window of a larger B (ldb > d) with beta == 0: the padding survives
    const int64_t ldb = d + 5;
    std::vector<double> Bp(ldb * n, -7.0);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S, (int64_t) 0, (int64_t) 0, A.data(), m, 0.0, Bp.data(), ldb);
    bool pad_ok = true, val_ok = true;
    for (int64_t j = 0; j < n; ++j) for (int64_t i = 0; i < ldb; ++i) {
        if (i >= d) { if (j < n - 1 || i < d) pad_ok = pad_ok && Bp[i + j * ldb] == -7.0; }
        else val_ok = val_ok && std::abs(Bp[i + j * ldb] - W[i + j * d]) <= 1e-12 * (1 + std::abs(W[i + j * d]));
    }
    CHECK(pad_ok && val_ok);
}

// device_pointers() mode: every array the caller passes already lives in device memory (rbx_malloc); the same calls give the
// host-mode results -- fill_dense, dense and sparse-operator sketch_general, left_spmm on a CSR / CSC matrix held on the device
// (used in place when compressed along the output axis, rbx_spmm_csx_*), sketch_sparse
template <typename T> static T* to_dev(const std::vector<T>& h) {
    void* p = nullptr;
    CHECK(rbx_malloc(&p, h.size() * sizeof(T)) == 0);
    CHECK(rbx_memcpy_h2d(p, h.data(), h.size() * sizeof(T), nullptr) == 0);
    return (T*) p;
}
template <typename T> static std::vector<T> to_host(const T* d, size_t n) {
    std::vector<T> h(n);
    CHECK(rbx_memcpy_d2h(h.data(), d, n * sizeof(T), nullptr) == 0);
    CHECK(rbx_stream_synchronize(nullptr) == 0);
    return h;
}
static void test_device_pointer_mode() {
    const int64_t d = 24, m = 700, n = 40;
    std::vector<double> A = randmat(m * n, 31), Bh(d * n, 0.0), B0(d * n, 0.0);
    DenseSkOp<double> Sd(DenseDist(d, m), RNGState<>(4));
    SparseSkOp<double> Ss({ d, m, 4, MajorAxis::Short }, RNGState<>(6));
    // host-mode results first
    std::vector<double> Wd(d * n, 0.0), Ws(d * n, 0.0), G(m * n);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, Sd, A.data(), m, 0.0, Wd.data(), d);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, Ss, A.data(), m, 0.0, Ws.data(), d);
    fill_dense(DenseDist(m, n), G.data(), RNGState<>(31));
    // a sparse data matrix (m x n, every third entry of a few columns) as CSR and CSC on the host
    std::vector<double> sv; std::vector<int64_t> sr, sc;
    for (int64_t i = 0; i < m; ++i) for (int64_t j = (i % 3); j < n; j += 5) { sr.push_back(i); sc.push_back(j); sv.push_back(std::cos(0.1 * i + j)); }
    const int64_t nnz = (int64_t) sv.size();
    std::vector<int64_t> rowptr(m + 1, 0), colptr(n + 1, 0);
    for (int64_t e = 0; e < nnz; ++e) { ++rowptr[sr[e] + 1]; ++colptr[sc[e] + 1]; }
    for (int64_t i = 0; i < m; ++i) rowptr[i + 1] += rowptr[i];
    for (int64_t j = 0; j < n; ++j) colptr[j + 1] += colptr[j];
    std::vector<int64_t> csc_rows(nnz); std::vector<double> csc_vals(nnz); std::vector<int64_t> cur(colptr.begin(), colptr.end() - 1);
    for (int64_t e = 0; e < nnz; ++e) { const int64_t p = cur[sc[e]]++; csc_rows[p] = sr[e]; csc_vals[p] = sv[e]; }
    // reference products on the host (dense loops): P1 = Asp^T (n x m) * X (m x 6), P2 = Asp (m x n) * Y (n x 6)
    const int64_t k = 6;
    std::vector<double> X = randmat(m * k, 77), Y = randmat(n * k, 78), P1(n * k, 0.0), P2(m * k, 0.0);
    for (int64_t e = 0; e < nnz; ++e) for (int64_t c = 0; c < k; ++c) { P1[sc[e] + c * n] += sv[e] * X[sr[e] + c * m]; P2[sr[e] + c * m] += sv[e] * Y[sc[e] + c * n]; }

    device_pointers() = true;
    double *dA = to_dev(A), *dB = to_dev(B0), *dG = to_dev(G), *dX = to_dev(X), *dY = to_dev(Y);
    double *dP1 = to_dev(P1), *dP2 = to_dev(P2);           // overwritten below (beta == 0)
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, Sd, dA, m, 0.0, dB, d);
    CHECK(max_rel(to_host(dB, d * n), Wd) < 1e-13);
    sketch_general(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, Ss, dA, m, 0.0, dB, d);
    CHECK(max_rel(to_host(dB, d * n), Ws) < 1e-13);
    fill_dense(DenseDist(m, n), dG, RNGState<>(31));
    CHECK(bits_equal(to_host(dG, m * n), G));
    {   // CSR and CSC views over device arrays (the containers do not own them)
        double *dv = to_dev(sv), *dcv = to_dev(csc_vals);
        int64_t *drp = to_dev(rowptr), *dci = to_dev(sc), *dcp = to_dev(colptr), *dri = to_dev(csc_rows);
        sparse_data::CSRMatrix<double> R(m, n, nnz, dv, drp, dci);
        sparse_data::CSCMatrix<double> C(m, n, nnz, dcv, dri, dcp);
        std::vector<double> got;
        left_spmm(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, m, k, n, 1.0, R, 0, 0, dY, n, 0.0, dP2, m);     // CSR in place
        got = to_host(dP2, m * k); CHECK(max_rel(got, P2) < 1e-13);
        left_spmm(blas::Layout::ColMajor, blas::Op::Trans, blas::Op::NoTrans, n, k, m, 1.0, C, 0, 0, dX, m, 0.0, dP1, n);       // CSC^T in place
        got = to_host(dP1, n * k); CHECK(max_rel(got, P1) < 1e-13);
        left_spmm(blas::Layout::ColMajor, blas::Op::Trans, blas::Op::NoTrans, n, k, m, 1.0, R, 0, 0, dX, m, 0.0, dP1, n);       // CSR^T: regrouped
        got = to_host(dP1, n * k); CHECK(max_rel(got, P1) < 1e-13);
        left_spmm(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, m, k, n, 1.0, C, 0, 0, dY, n, 0.0, dP2, m);     // CSC: regrouped
        got = to_host(dP2, m * k); CHECK(max_rel(got, P2) < 1e-13);
        // sketch_sparse with the matrix on the device against the host-mode call
        DenseSkOp<double> S2(DenseDist(d, m), RNGState<>(8));
        std::vector<double> Wsp(d * n, 0.0);
        device_pointers() = false;
        sparse_data::CSRMatrix<double> Rh(m, n, nnz, sv.data(), rowptr.data(), sc.data());
        sketch_sparse(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S2, 0, 0, Rh, 0, 0, 0.0, Wsp.data(), d);
        device_pointers() = true;
        sketch_sparse(blas::Layout::ColMajor, blas::Op::NoTrans, blas::Op::NoTrans, d, n, m, 1.0, S2, 0, 0, R, 0, 0, 0.0, dB, d);
        CHECK(max_rel(to_host(dB, d * n), Wsp) < 1e-13);
        rbx_free(dv); rbx_free(dcv); rbx_free(drp); rbx_free(dci); rbx_free(dcp); rbx_free(dri);
    }
    device_pointers() = false;
    rbx_free(dA); rbx_free(dB); rbx_free(dG); rbx_free(dX); rbx_free(dY); rbx_free(dP1); rbx_free(dP2);
}

int main() {
    if (rbx_device_count() < 1) { std::printf("no CUDA device: the product path has no CPU fallback\n"); return 77; }
    test_fill_dense();
    test_fill_sparse();
    test_sketch_dense();
    test_sketch_sparseop();
    test_sketch_sparse();
    test_index_sampling();
    test_rngstate_surface();
    test_sparse_surface();
    test_device_pointer_mode();
    std::printf(failures ? "FAILED (%d)\n" : "OK\n", failures);
    return failures ? 1 : 0;
}
