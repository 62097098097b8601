// ORACLE / TEST INFRASTRUCTURE ONLY -- never part of the shipped product path.
//
// Stand-in for BLAS++ <blas.hh> (icl-utk-edu/blaspp, un-pinned by the reference:
// /root/reference/.github/workflows/core-linux.yaml:26-33). BLAS++ is not vendored in
// /root/reference and not installed here, so the reference headers cannot compile
// without this. Only the subset the reference library proper touches is provided:
//   enums Layout/Op/Uplo/Side/Diag        (base.hh:70-114, skge.hh, sksy.hh, ...)
//   gemm   (skge.hh:213,362)               scal (dense_skops.hh:512, coo_spmm_impl.hh:120, util.hh:57,268)
//   axpy   (csc_spmm_impl.hh:108,205; csr_spmm_impl.hh:87,151)     copy (dense_skops.hh:525)
// plus gemv/symm/dot/nrm2 used by the oracle-side checks.
//
// gemm/axpy/scal forward to the multithreaded OpenBLAS that scipy bundles (symbols
// scipy_cblas_*; found with dlopen at run time) when it can be located, else to
// plain OpenMP loops. `blas_shim::backend_name()` says which one is live.
#pragma once

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <complex>
#include <string>
#include <vector>
#include <limits>
#include <algorithm>
#include <stdexcept>
#include <dlfcn.h>
#include <glob.h>

namespace blas {

enum class Layout : char { ColMajor = 'C', RowMajor = 'R' };
enum class Op     : char { NoTrans  = 'N', Trans    = 'T', ConjTrans = 'C' };
enum class Uplo   : char { Upper    = 'U', Lower    = 'L', General   = 'G' };
enum class Diag   : char { NonUnit  = 'N', Unit     = 'U' };
enum class Side   : char { Left     = 'L', Right    = 'R' };

}  // namespace blas

namespace blas_shim {

// CBLAS enum values (fixed by the CBLAS standard).
enum { CblasRowMajor = 101, CblasColMajor = 102, CblasNoTrans = 111, CblasTrans = 112 };

typedef void (*sgemm_t)(int, int, int, int, int, int, float, const float*, int, const float*, int, float, float*, int);
typedef void (*dgemm_t)(int, int, int, int, int, int, double, const double*, int, const double*, int, double, double*, int);
typedef void (*saxpy_t)(int, float, const float*, int, float*, int);
typedef void (*daxpy_t)(int, double, const double*, int, double*, int);
typedef void (*sscal_t)(int, float, float*, int);
typedef void (*dscal_t)(int, double, double*, int);
typedef void (*setthr_t)(int);
typedef int  (*getthr_t)(void);

struct Backend {
    void*   handle = nullptr;
    sgemm_t sgemm = nullptr;  dgemm_t dgemm = nullptr;
    saxpy_t saxpy = nullptr;  daxpy_t daxpy = nullptr;
    sscal_t sscal = nullptr;  dscal_t dscal = nullptr;
    setthr_t set_threads = nullptr; getthr_t get_threads = nullptr;
    std::string name = "openmp-loops";
};

inline Backend& backend() {
    static Backend be = []() {
        Backend b;
        if (getenv("RBX_ORACLE_NO_OPENBLAS")) return b;
        std::vector<std::string> candidates;
        if (const char* p = getenv("RBX_ORACLE_OPENBLAS")) candidates.push_back(p);
        const char* pats[] = {
            "/opt/prime-rl/.venv/lib/python3*/site-packages/scipy.libs/libscipy_openblas-*.so",
            "/usr/lib/python3*/site-packages/scipy.libs/libscipy_openblas-*.so",
        };
        for (const char* pat : pats) {
            glob_t g; memset(&g, 0, sizeof g);
            if (glob(pat, 0, nullptr, &g) == 0)
                for (size_t i = 0; i < g.gl_pathc; ++i) candidates.push_back(g.gl_pathv[i]);
            globfree(&g);
        }
        for (auto& c : candidates) {
            void* h = dlopen(c.c_str(), RTLD_NOW | RTLD_LOCAL);
            if (!h) continue;
            b.sgemm = (sgemm_t) dlsym(h, "scipy_cblas_sgemm");
            b.dgemm = (dgemm_t) dlsym(h, "scipy_cblas_dgemm");
            b.saxpy = (saxpy_t) dlsym(h, "scipy_cblas_saxpy");
            b.daxpy = (daxpy_t) dlsym(h, "scipy_cblas_daxpy");
            b.sscal = (sscal_t) dlsym(h, "scipy_cblas_sscal");
            b.dscal = (dscal_t) dlsym(h, "scipy_cblas_dscal");
            b.set_threads = (setthr_t) dlsym(h, "scipy_openblas_set_num_threads");
            b.get_threads = (getthr_t) dlsym(h, "scipy_openblas_get_num_threads");
            if (b.sgemm && b.dgemm && b.saxpy && b.daxpy && b.sscal && b.dscal) {
                b.handle = h; b.name = "scipy-openblas:" + c; return b;
            }
            dlclose(h);
            b = Backend();
        }
        return b;
    }();
    return be;
}

inline const char* backend_name() { return backend().name.c_str(); }

template <typename T>
inline T at(const T* A, int64_t ld, bool colmajor, bool trans, int64_t i, int64_t j) {
    // element (i,j) of op(A)
    if (trans) std::swap(i, j);
    return colmajor ? A[i + j * ld] : A[i * ld + j];
}

template <typename T>
void gemm_loops(bool colmajor, bool ta, bool tb, int64_t m, int64_t n, int64_t k,
                T alpha, const T* A, int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
    #pragma omp parallel for collapse(2) schedule(static)
    for (int64_t i = 0; i < m; ++i) {
        for (int64_t j = 0; j < n; ++j) {
            T acc = 0;
            for (int64_t l = 0; l < k; ++l)
                acc += at(A, lda, colmajor, ta, i, l) * at(B, ldb, colmajor, tb, l, j);
            T* c = colmajor ? &C[i + j * ldc] : &C[i * ldc + j];
            *c = (beta == (T) 0) ? alpha * acc : alpha * acc + beta * (*c);
        }
    }
}

}  // namespace blas_shim

namespace blas {

inline void require_int(int64_t v) {
    if (v > (int64_t) std::numeric_limits<int>::max() || v < (int64_t) std::numeric_limits<int>::min())
        throw std::overflow_error("blas shim: dimension exceeds 32-bit CBLAS int");
}

template <typename T>
void gemm(Layout layout, Op transA, Op transB, int64_t m, int64_t n, int64_t k,
          T alpha, const T* A, int64_t lda, const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
    bool cm = layout == Layout::ColMajor;
    bool ta = transA != Op::NoTrans, tb = transB != Op::NoTrans;
    if (m == 0 || n == 0) return;
    auto& be = blas_shim::backend();
    if (be.handle && k > 0) {
        require_int(m); require_int(n); require_int(k); require_int(lda); require_int(ldb); require_int(ldc);
        int lay = cm ? blas_shim::CblasColMajor : blas_shim::CblasRowMajor;
        int ota = ta ? blas_shim::CblasTrans : blas_shim::CblasNoTrans;
        int otb = tb ? blas_shim::CblasTrans : blas_shim::CblasNoTrans;
        if constexpr (std::is_same_v<T, float>)
            be.sgemm(lay, ota, otb, (int) m, (int) n, (int) k, alpha, A, (int) lda, B, (int) ldb, beta, C, (int) ldc);
        else
            be.dgemm(lay, ota, otb, (int) m, (int) n, (int) k, alpha, A, (int) lda, B, (int) ldb, beta, C, (int) ldc);
        return;
    }
    blas_shim::gemm_loops(cm, ta, tb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
}

template <typename T>
void scal(int64_t n, T alpha, T* x, int64_t incx) {
    auto& be = blas_shim::backend();
    if (be.handle && n <= std::numeric_limits<int>::max() && std::abs(incx) <= std::numeric_limits<int>::max()) {
        if constexpr (std::is_same_v<T, float>) be.sscal((int) n, alpha, x, (int) incx);
        else be.dscal((int) n, alpha, x, (int) incx);
        return;
    }
    #pragma omp parallel for schedule(static) if (n > 100000)
    for (int64_t i = 0; i < n; ++i) x[i * incx] *= alpha;
}

template <typename T>
void axpy(int64_t n, T alpha, const T* x, int64_t incx, T* y, int64_t incy) {
    auto& be = blas_shim::backend();
    // small n: avoid library-call + threading overhead inside the reference's own omp regions
    if (be.handle && n >= 256 && n <= std::numeric_limits<int>::max()) {
        if constexpr (std::is_same_v<T, float>) be.saxpy((int) n, alpha, x, (int) incx, y, (int) incy);
        else be.daxpy((int) n, alpha, x, (int) incx, y, (int) incy);
        return;
    }
    for (int64_t i = 0; i < n; ++i) y[i * incy] += alpha * x[i * incx];
}

template <typename T>
void copy(int64_t n, const T* x, int64_t incx, T* y, int64_t incy) {
    if (incx == 1 && incy == 1) { std::memcpy(y, x, (size_t) n * sizeof(T)); return; }
    for (int64_t i = 0; i < n; ++i) y[i * incy] = x[i * incx];
}

template <typename T>
void gemv(Layout layout, Op trans, int64_t m, int64_t n, T alpha, const T* A, int64_t lda,
          const T* x, int64_t incx, T beta, T* y, int64_t incy) {
    bool cm = layout == Layout::ColMajor;
    bool t = trans != Op::NoTrans;
    int64_t leny = t ? n : m, lenx = t ? m : n;
    for (int64_t i = 0; i < leny; ++i) {
        T acc = 0;
        for (int64_t j = 0; j < lenx; ++j) acc += blas_shim::at(A, lda, cm, t, i, j) * x[j * incx];
        y[i * incy] = (beta == (T) 0) ? alpha * acc : alpha * acc + beta * y[i * incy];
    }
}

template <typename T>
void symm(Layout layout, Side side, Uplo uplo, int64_t m, int64_t n, T alpha, const T* A, int64_t lda,
          const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {
    bool cm = layout == Layout::ColMajor;
    int64_t ka = (side == Side::Left) ? m : n;
    auto sym = [&](int64_t i, int64_t j) {
        bool upper_entry = i <= j;
        bool stored = (uplo == Uplo::Upper) ? upper_entry : !upper_entry || i == j;
        if (!stored) std::swap(i, j);
        return cm ? A[i + j * lda] : A[i * lda + j];
    };
    (void) ka;
    for (int64_t i = 0; i < m; ++i) for (int64_t j = 0; j < n; ++j) {
        T acc = 0;
        if (side == Side::Left) for (int64_t l = 0; l < m; ++l) acc += sym(i, l) * (cm ? B[l + j * ldb] : B[l * ldb + j]);
        else                    for (int64_t l = 0; l < n; ++l) acc += (cm ? B[i + l * ldb] : B[i * ldb + l]) * sym(l, j);
        T* c = cm ? &C[i + j * ldc] : &C[i * ldc + j];
        *c = (beta == (T) 0) ? alpha * acc : alpha * acc + beta * (*c);
    }
}

template <typename T>
T dot(int64_t n, const T* x, int64_t incx, const T* y, int64_t incy) {
    T acc = 0;
    for (int64_t i = 0; i < n; ++i) acc += x[i * incx] * y[i * incy];
    return acc;
}

template <typename T>
T nrm2(int64_t n, const T* x, int64_t incx) {
    long double acc = 0;
    for (int64_t i = 0; i < n; ++i) acc += (long double) x[i * incx] * x[i * incx];
    return (T) std::sqrt((double) acc);
}

}  // namespace blas
