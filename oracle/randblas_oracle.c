/* ORACLE / TEST INFRASTRUCTURE ONLY -- never part of the shipped product path.
 *
 * Plain-C CPU restatement of the RandBLAS sketching hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 * Each function cites the reference file:line it follows (paths under /root/reference).
 *
 * Pinning (see tests/test_oracle_pins.py):
 *   - Philox4x32-10 and counter carries: pinned by the reference's KATs
 *     (test/test_basic_rng/r123_kat_vectors.txt:19-21, test_r123.cc:679-766).
 *   - everything else: pinned against outputs of the reference itself compiled here
 *     (oracle/_ref/librandblas_ref.so, built from the untouched reference headers against
 *     oracle/shims) and against fixtures under tests/golden/ generated from it.
 *   - the float transforms u01/uneg11/boxmuller restate Random123's public definitions;
 *     a real Random123 checkout is not available offline => that one link is PARITY UNPINNED.
 *
 * Built with -ffp-contract=off so that the only fused multiply-adds are the explicit fma()
 * calls in the libm models below.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { uint32_t c[4]; uint32_t k[2]; } orc_state;

/* ------------------------------------------------------------------------------------------
 * Philox4x32-10. Random123 philox.h (not in tree); call sites random_gen.hh:143,171,
 * sparse_skops.hh:78. KATs: test/test_basic_rng/r123_kat_vectors.txt:19-21.
 * ---------------------------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t) 0xD2511F53u * c0;
        uint64_t p1 = (uint64_t) 0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t) p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t) p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* 128-bit little-endian counter += n. Random123 array.h incr(); pinned by test_r123.cc:679-766. */
void orc_ctr_add(uint32_t ctr[4], uint64_t n) {
    uint64_t lo = ((uint64_t) ctr[1] << 32) | ctr[0];
    uint64_t hi = ((uint64_t) ctr[3] << 32) | ctr[2];
    uint64_t s = lo + n;
    if (s < lo) hi += 1;
    ctr[0] = (uint32_t) s; ctr[1] = (uint32_t)(s >> 32);
    ctr[2] = (uint32_t) hi; ctr[3] = (uint32_t)(hi >> 32);
}

/* Random123 uniform.hpp: uneg11<float>(u32), u01<float>(u32). Call site random_gen.hh:171. */
float orc_uneg11(uint32_t u) { return (float)(int32_t) u * 0x1p-31f + 0x1p-32f; }
float orc_u01(uint32_t u)    { return (float) u * 0x1p-32f + 0x1p-33f; }

/* Random123 boxmuller.hpp (float) with the reference's host sincospif (random_gen.hh:61-65):
 * sincosf(PIf * x); r = sqrtf(-2.f * logf(u)). Uses the host libm, like the reference. */
void orc_boxmuller(uint32_t u0, uint32_t u1, float* s_out, float* c_out) {
    const float PIf = 3.1415926535897932f;
    float s, c;
    sincosf(PIf * orc_uneg11(u0), &s, &c);
    float r = sqrtf(-2.f * logf(orc_u01(u1)));
    *s_out = s * r;
    *c_out = c * r;
}

/* r123ext::boxmul::generate / r123ext::uneg11::generate (random_gen.hh:96-173): 4 floats per counter. */
void orc_generate4(char family, const uint32_t ctr[4], const uint32_t key[2], float out[4]) {
    uint32_t r[4];
    orc_philox4x32_10(ctr, key, r);
    if (family == 'G') {
        orc_boxmuller(r[0], r[1], &out[0], &out[1]);
        orc_boxmuller(r[2], r[3], &out[2], &out[3]);
    } else {
        for (int i = 0; i < 4; ++i) out[i] = orc_uneg11(r[i]);
    }
}

/* ------------------------------------------------------------------------------------------
 * Models of glibc 2.39's x86-64 FMA variants of sincosf and logf (the IFUNC targets on any
 * FMA-capable host). These are what the CUDA kernel implements in FP64 with explicit
 * __fma_rn; the operation order and fusion pattern were read from the disassembly of
 * libm.so.6 (DESIGN.md "bit-exact Gaussian"). orc_sweep_* below compares them with the
 * live libm over the full 2^32 input domains.
 * ---------------------------------------------------------------------------------------- */
static const double SC_HPI_INV = 0x1.45f306dc9c883p+23;  /* 2/pi * 2^24 */
static const double SC_HPI     = 0x1.921fb54442d18p+0;
static const double SC_C0 = 0x1p0, SC_C1 = -0x1.ffffffd0c621cp-2, SC_C2 = 0x1.55553e1068f19p-5,
                    SC_C3 = -0x1.6c087e89a359dp-10, SC_C4 = 0x1.99343027bf8c3p-16;
static const double SC_S1 = -0x1.555545995a603p-3, SC_S2 = 0x1.1107605230bc4p-7,
                    SC_S3 = -0x1.994eb3774cf24p-13;

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

void orc_sincosf_model(float y, float* sinp, float* cosp) {
    uint32_t top = (f2u(y) >> 20) & 0x7ff;
    double x = (double) y;
    if (top < 0x3f4) {                       /* |y| < ~pi/4 */
        if (top < 0x398) { *sinp = y; *cosp = 1.0f; return; }   /* |y| < 2^-12 */
        double x2 = x * x;
        double x3 = x * x2, x4 = x2 * x2;
        double s1 = fma(x2, SC_S3, SC_S2), c2 = fma(x2, SC_C4, SC_C3);
        double c1 = fma(x2, SC_C1, SC_C0);
        double x5 = x2 * x3, x6 = x2 * x4;
        double s = fma(x3, SC_S1, x), c = fma(x4, SC_C2, c1);
        *sinp = (float) fma(s1, x5, s);
        *cosp = (float) fma(c2, x6, c);
        return;
    }
    /* pi/4 <= |y| < 120 : reduce_fast without TOINT_INTRINSICS */
    double r = x * SC_HPI_INV;
    int n = ((int32_t) r + 0x800000) >> 24;
    double xr = fma(-(double) n, SC_HPI, x);
    double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;   /* sign[] = {1,-1,-1,1} */
    double flip = (n & 2) ? -1.0 : 1.0;                          /* table[1] negates the cosine poly */
    double xs = xr * sign, x2 = xr * xr;
    double s1 = fma(x2, SC_S3, SC_S2), c2 = fma(x2, flip * SC_C4, flip * SC_C3);
    double x3 = x2 * xs, x4 = x2 * x2;
    double x5 = x2 * x3, x6 = x2 * x4;
    double c1 = fma(x2, flip * SC_C1, flip * SC_C0);
    double s = fma(x3, SC_S1, xs), c = fma(x4, flip * SC_C2, c1);
    float ps = (float) fma(s1, x5, s), pc = (float) fma(c2, x6, c);
    if (n & 1) { *cosp = ps; *sinp = pc; } else { *sinp = ps; *cosp = pc; }
}

static const double LG_INVC[16] = {
    0x1.661ec79f8f3bep+0, 0x1.571ed4aaf883dp+0, 0x1.49539f0f010b0p+0, 0x1.3c995b0b80385p+0,
    0x1.30d190c8864a5p+0, 0x1.25e227b0b8ea0p+0, 0x1.1bb4a4a1a343fp+0, 0x1.12358f08ae5bap+0,
    0x1.0953f419900a7p+0, 0x1.0000000000000p+0, 0x1.e608cfd9a47acp-1, 0x1.ca4b31f026aa0p-1,
    0x1.b2036576afce6p-1, 0x1.9c2d163a1aa2dp-1, 0x1.886e6037841edp-1, 0x1.767dcf5534862p-1 };
static const double LG_LOGC[16] = {
    -0x1.57bf7808caadep-2, -0x1.2bef0a7c06ddbp-2, -0x1.01eae7f513a67p-2, -0x1.b31d8a68224e9p-3,
    -0x1.6574f0ac07758p-3, -0x1.1aa2bc79c8100p-3, -0x1.a4e76ce8c0e5ep-4, -0x1.1973c5a611cccp-4,
    -0x1.252f438e10c1ep-5, 0x0p+0, 0x1.aa5aa5df25984p-5, 0x1.c5e53aa362eb4p-4,
    0x1.526e57720db08p-3, 0x1.bc2860d224770p-3, 0x1.1058bc8a07ee1p-2, 0x1.4043057b6ee09p-2 };
static const double LG_LN2 = 0x1.62e42fefa39efp-1;
static const double LG_A0 = -0x1.00ea348b88334p-2, LG_A1 = 0x1.5575b0be00b6ap-2, LG_A2 = -0x1.ffffef20a4123p-2;

/* valid for positive normal x (the only inputs u01<float> can produce) */
float orc_logf_model(float xf) {
    uint32_t ix = f2u(xf);
    if (ix == 0x3f800000u) return 0.0f;
    uint32_t tmp = ix - 0x3f330000u;
    int i = (tmp >> 19) & 15;
    int k = (int32_t) tmp >> 23;
    uint32_t iz = ix - (tmp & 0xff800000u);
    double z = (double) u2f(iz);
    double r = fma(z, LG_INVC[i], -1.0);
    double y0 = fma((double) k, LG_LN2, LG_LOGC[i]);
    double r2 = r * r;
    double y = fma(r, LG_A1, LG_A2);
    y = fma(r2, LG_A0, y);
    y = fma(y, r2, y0 + r);
    return (float) y;
}

void orc_boxmuller_model(uint32_t u0, uint32_t u1, float* s_out, float* c_out) {
    const float PIf = 3.1415926535897932f;
    float s, c;
    orc_sincosf_model(PIf * orc_uneg11(u0), &s, &c);
    float r = sqrtf(-2.f * orc_logf_model(orc_u01(u1)));
    *s_out = s * r;
    *c_out = c * r;
}

/* boxmuller through the live host libm (model == 0) or the explicit-FMA models (model == 1) on arrays */
void orc_boxmuller_arr(int64_t n, const uint32_t* u0, const uint32_t* u1, float* s, float* c, int model) {
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        if (model) orc_boxmuller_model(u0[i], u1[i], &s[i], &c[i]);
        else orc_boxmuller(u0[i], u1[i], &s[i], &c[i]);
    }
}

/* Compare the models with the live libm over u in [lo, hi) with the given stride.
 * Returns the number of inputs whose float bits differ; first_bad gets the first such u. */
int64_t orc_sweep_sincos(uint64_t lo, uint64_t hi, uint64_t stride, uint32_t* first_bad) {
    const float PIf = 3.1415926535897932f;
    int64_t bad = 0;
    uint64_t fb = UINT64_MAX;
    #pragma omp parallel for schedule(static) reduction(+:bad) reduction(min:fb)
    for (uint64_t t = 0; t < (hi - lo + stride - 1) / stride; ++t) {
        uint64_t u = lo + t * stride;
        float y = PIf * orc_uneg11((uint32_t) u);
        float s0, c0, s1, c1;
        sincosf(y, &s0, &c0);
        orc_sincosf_model(y, &s1, &c1);
        if (f2u(s0) != f2u(s1) || f2u(c0) != f2u(c1)) { bad++; if (u < fb) fb = u; }
    }
    if (first_bad && fb != UINT64_MAX) *first_bad = (uint32_t) fb;
    return bad;
}
int64_t orc_sweep_log(uint64_t lo, uint64_t hi, uint64_t stride, uint32_t* first_bad) {
    int64_t bad = 0;
    uint64_t fb = UINT64_MAX;
    #pragma omp parallel for schedule(static) reduction(+:bad) reduction(min:fb)
    for (uint64_t t = 0; t < (hi - lo + stride - 1) / stride; ++t) {
        uint64_t u = lo + t * stride;
        float x = orc_u01((uint32_t) u);
        if (f2u(logf(x)) != f2u(orc_logf_model(x))) { bad++; if (u < fb) fb = u; }
    }
    if (first_bad && fb != UINT64_MAX) *first_bad = (uint32_t) fb;
    return bad;
}

/* ------------------------------------------------------------------------------------------
 * Dense operators.
 * ---------------------------------------------------------------------------------------- */

/* One entry of the implicit row-major parent matrix with n_cols columns, rows padded to a
 * multiple of 4 so each row starts a fresh counter: dense_skops.hh:109-131 (pad, ctr_mat_start,
 * ctr_inter_row_stride) and :136-162 (row loop). Entry (row, col) of the PARENT is lane col%4 of
 * OP(philox(seed.ctr + row*W + col/4, key)), W = ceil(n_cols/4). */
static float parent_entry(char family, int64_t n_cols, int64_t row, int64_t col, const orc_state* st) {
    int64_t W = (n_cols + 3) / 4;
    uint32_t c[4] = { st->c[0], st->c[1], st->c[2], st->c[3] };
    orc_ctr_add(c, (uint64_t)(row * W + col / 4));
    float v[4];
    orc_generate4(family, c, st->k, v);
    return v[col % 4];
}

/* fill_dense_submat_impl (dense_skops.hh:96-170): smat[i*lda + j] = parent(r0+i, c0+j), with
 * r0 = ptr / n_cols, c0 = ptr % n_cols. next = seed.ctr + ptr_padded/4 + n_srows*W (:166-169). */
#define DEF_SUBMAT(SUF, T)                                                                              \
int orc_fill_dense_submat_##SUF(char family, int64_t n_cols, T* smat, int64_t n_srows, int64_t n_scols,  \
                                int64_t ptr, const uint32_t ctr[4], const uint32_t key[2], int64_t lda,   \
                                uint32_t next_ctr[4]) {                                                   \
    if (lda <= 0) lda = n_scols; else if (lda < n_scols) return 1;                                        \
    if (n_cols < n_scols) return 1;                                                                       \
    orc_state st; memcpy(st.c, ctr, 16); memcpy(st.k, key, 8);                                            \
    int64_t r0 = ptr / n_cols, c0 = ptr % n_cols, W = (n_cols + 3) / 4;                                   \
    _Pragma("omp parallel for schedule(static)")                                                          \
    for (int64_t i = 0; i < n_srows; ++i) {                                                               \
        int64_t j = 0;                                                                                    \
        while (j < n_scols) {                                                                             \
            int64_t col = c0 + j, blk = col / 4;                                                          \
            uint32_t c[4] = { st.c[0], st.c[1], st.c[2], st.c[3] };                                       \
            orc_ctr_add(c, (uint64_t)((r0 + i) * W + blk));                                               \
            float v[4];                                                                                   \
            orc_generate4(family, c, st.k, v);                                                            \
            for (int l = (int)(col % 4); l < 4 && j < n_scols; ++l, ++j) smat[i * lda + j] = (T) v[l];    \
        }                                                                                                 \
    }                                                                                                     \
    if (next_ctr) {                                                                                       \
        memcpy(next_ctr, ctr, 16);                                                                        \
        orc_ctr_add(next_ctr, (uint64_t)(r0 * W + c0 / 4));                                               \
        orc_ctr_add(next_ctr, (uint64_t)(n_srows * W));                                                   \
    }                                                                                                     \
    return 0;                                                                                             \
}
DEF_SUBMAT(f32, float)
DEF_SUBMAT(f64, double)

/* dist_to_layout / major_axis_length (dense_skops.hh:297-316). */
char orc_dist_to_layout(int64_t n_rows, int64_t n_cols, char major_axis) {
    int wide = n_rows < n_cols, lng = major_axis == 'L';
    if (wide && lng) return 'R';
    if (wide) return 'C';
    if (lng) return 'C';
    return 'R';
}
int64_t orc_major_axis_length(int64_t n_rows, int64_t n_cols, char major_axis) {
    int64_t mx = n_rows > n_cols ? n_rows : n_cols, mn = n_rows < n_cols ? n_rows : n_cols;
    return major_axis == 'L' ? mx : mn;
}
/* dense::compute_next_state (dense_skops.hh:172-191). */
void orc_dense_next_state(int64_t n_rows, int64_t n_cols, char major_axis, const uint32_t ctr[4], uint32_t out[4]) {
    int64_t major = orc_major_axis_length(n_rows, n_cols, major_axis);
    int64_t minor = n_rows + (n_cols - major);
    memcpy(out, ctr, 16);
    orc_ctr_add(out, (uint64_t)(((major + 3) / 4) * minor));
}

/* Entry (i, j) of the full implicit operator S ~ D (before any layout question):
 * natural RowMajor => parent is S itself with ma_len columns; natural ColMajor => parent is S^T
 * (dense_skops.hh:491-503). Uniform entries are post-scaled by (T)sqrt(3) (:511-513). */
static float dist_entry_raw(int64_t D_rows, int64_t D_cols, char family, char major_axis,
                            const orc_state* st, int64_t i, int64_t j) {
    int64_t ma = orc_major_axis_length(D_rows, D_cols, major_axis);
    if (orc_dist_to_layout(D_rows, D_cols, major_axis) == 'R') return parent_entry(family, ma, i, j, st);
    return parent_entry(family, ma, j, i, st);
}

/* fill_dense(layout, D, n_rows, n_cols, ro_s, co_s, buff, seed) (dense_skops.hh:486-532):
 * mat(buff)[i,j] = S[ro_s+i, co_s+j], stored compactly in `layout`. */
#define DEF_FILL(SUF, T)                                                                                \
int orc_fill_dense_##SUF(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,      \
                         int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, T* buff,             \
                         const uint32_t ctr[4], const uint32_t key[2], uint32_t next_ctr[4]) {            \
    if (D_rows < n_rows + ro_s || D_cols < n_cols + co_s) return 1;                                       \
    if (family != 'G' && family != 'U') return 2;                                                         \
    orc_state st; memcpy(st.c, ctr, 16); memcpy(st.k, key, 8);                                            \
    const T scale = (T) sqrt(3.0);                                                                        \
    _Pragma("omp parallel for schedule(static)")                                                          \
    for (int64_t i = 0; i < n_rows; ++i)                                                                  \
        for (int64_t j = 0; j < n_cols; ++j) {                                                            \
            T v = (T) dist_entry_raw(D_rows, D_cols, family, major_axis, &st, ro_s + i, co_s + j);        \
            if (family == 'U') v = v * scale;                                                             \
            if (layout == 'C') buff[i + j * n_rows] = v; else buff[i * n_cols + j] = v;                   \
        }                                                                                                 \
    if (next_ctr) {                                                                                       \
        /* return value of fill_dense_submat_impl for the natural-layout call (:494-511) */               \
        int64_t ma = orc_major_axis_length(D_rows, D_cols, major_axis), W = (ma + 3) / 4;                 \
        int nat_c = orc_dist_to_layout(D_rows, D_cols, major_axis) == 'C';                                \
        int64_t ptr = nat_c ? ro_s + co_s * ma : ro_s * ma + co_s;                                        \
        int64_t nsr = nat_c ? n_cols : n_rows;                                                            \
        memcpy(next_ctr, ctr, 16);                                                                        \
        orc_ctr_add(next_ctr, (uint64_t)((ptr / ma) * W + (ptr % ma) / 4));                               \
        orc_ctr_add(next_ctr, (uint64_t)(nsr * W));                                                       \
    }                                                                                                     \
    return 0;                                                                                             \
}
DEF_FILL(f32, float)
DEF_FILL(f64, double)

/* ------------------------------------------------------------------------------------------
 * Sparse operators: repeated Fisher-Yates (sparse_skops.hh:53-106), restated with a lazy
 * permutation (only displaced positions are remembered), which is also what the CUDA kernel does.
 * For vector i and step j: rv = philox(seed.ctr + i*vec_nnz + j); ell = j + rv[0] % (dim_major - j);
 * emit work[ell]; then work[ell] = work[j]. Every vector restarts from the identity (:93-103).
 * ---------------------------------------------------------------------------------------- */
#define DEF_FY(SUF, T, I)                                                                               \
int orc_repeated_fisher_yates_##SUF(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz,       \
                                    int64_t dim_major, int64_t dim_minor, I* idxs_major, I* idxs_minor,   \
                                    T* vals) {                                                            \
    if (vec_nnz > dim_major) return 1;                                                                    \
    _Pragma("omp parallel")                                                                               \
    {                                                                                                     \
        int64_t* pos = (int64_t*) malloc(sizeof(int64_t) * (size_t) vec_nnz);                             \
        int64_t* val = (int64_t*) malloc(sizeof(int64_t) * (size_t) vec_nnz);                             \
        _Pragma("omp for schedule(static)")                                                               \
        for (int64_t i = 0; i < dim_minor; ++i) {                                                         \
            int64_t nov = 0;                                                                              \
            for (int64_t j = 0; j < vec_nnz; ++j) {                                                       \
                uint32_t c[4] = { ctr[0], ctr[1], ctr[2], ctr[3] }, rv[4];                                \
                orc_ctr_add(c, (uint64_t)(i * vec_nnz + j));                                              \
                orc_philox4x32_10(c, key, rv);                                                            \
                int64_t ell = j + (int64_t)(rv[0] % (uint64_t)(dim_major - j));                           \
                int64_t at_ell = ell, at_j = j, slot = -1;                                                \
                for (int64_t t = 0; t < nov; ++t) {                                                       \
                    if (pos[t] == ell) { at_ell = val[t]; slot = t; }                                     \
                    if (pos[t] == j) at_j = val[t];                                                       \
                }                                                                                         \
                if (ell != j) {                                                                           \
                    if (slot < 0) { slot = nov++; pos[slot] = ell; }                                      \
                    val[slot] = at_j;                                                                     \
                }                                                                                         \
                idxs_major[i * vec_nnz + j] = (I) at_ell;                                                 \
                if (vals) vals[i * vec_nnz + j] = (rv[1] % 2 == 0) ? (T) 1.0 : (T) -1.0;                  \
                if (idxs_minor) idxs_minor[i * vec_nnz + j] = (I) i;                                      \
            }                                                                                             \
        }                                                                                                 \
        free(pos); free(val);                                                                             \
    }                                                                                                     \
    return 0;                                                                                             \
}
DEF_FY(f32_i32, float, int32_t)
DEF_FY(f32_i64, float, int64_t)
DEF_FY(f64_i32, double, int32_t)
DEF_FY(f64_i64, double, int64_t)

/* fill_sparse (sparse_skops.hh:389-413) + sparse::compute_next_state (:115-126, reproduced as-is:
 * SASO advances by vec_nnz*min(dims), LASO by vec_nnz*max(dims)). */
#define DEF_FILL_SPARSE(SUF, T, I)                                                                      \
int orc_fill_sparse_##SUF(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis,              \
                          const uint32_t ctr[4], const uint32_t key[2], I* rows, I* cols, T* vals,        \
                          uint32_t next_ctr[4]) {                                                         \
    int64_t lng = n_rows > n_cols ? n_rows : n_cols, sht = n_rows < n_cols ? n_rows : n_cols;             \
    int is_wide = n_rows == sht;                                                                          \
    I* short_idx = is_wide ? rows : cols;                                                                 \
    I* long_idx = is_wide ? cols : rows;                                                                  \
    int rc = (major_axis == 'S')                                                                          \
        ? orc_repeated_fisher_yates_##SUF(ctr, key, vec_nnz, sht, lng, short_idx, long_idx, vals)         \
        : orc_repeated_fisher_yates_##SUF(ctr, key, vec_nnz, lng, sht, long_idx, short_idx, vals);        \
    if (next_ctr) {                                                                                       \
        memcpy(next_ctr, ctr, 16);                                                                        \
        orc_ctr_add(next_ctr, (uint64_t)(vec_nnz * (major_axis == 'S' ? sht : lng)));                     \
    }                                                                                                     \
    return rc;                                                                                            \
}
DEF_FILL_SPARSE(f32_i32, float, int32_t)
DEF_FILL_SPARSE(f32_i64, float, int64_t)
DEF_FILL_SPARSE(f64_i32, double, int32_t)
DEF_FILL_SPARSE(f64_i64, double, int64_t)

/* ------------------------------------------------------------------------------------------
 * Sketching: definitions by explicit operators (double accumulation). These state WHAT
 * lskge3/rskge3 (skge.hh:173-364), lskges/rskges (skge.hh:485-641) and lsksp3/rsksp3
 * (sksp.hh:147-345) compute: B = alpha*op(submat(S))*op(A) + beta*B (left) or
 * B = alpha*op(A)*op(submat(S)) + beta*B (right), submat(S) = S[ro_s:, co_s:]. Results are
 * compared with the reference's tolerance (test/test_matmul_cores/linop_common.hh:257-263).
 * ---------------------------------------------------------------------------------------- */

/* C(MxN) = alpha * P(MxK) * Q(KxN) + beta*C, P and Q row-major double, C in `layout` with ldc.
 * beta == 0 => C is not read (BLAS convention; util.hh:51-59 safe_scal). */
#define DEF_GEMM_DEF(SUF, T)                                                                            \
static void gemm_def_##SUF(char layout, int64_t M, int64_t N, int64_t K, double alpha,                   \
                           const double* P, const double* Q, double beta, T* C, int64_t ldc) {            \
    _Pragma("omp parallel for schedule(static)")                                                          \
    for (int64_t i = 0; i < M; ++i)                                                                       \
        for (int64_t j = 0; j < N; ++j) {                                                                 \
            double acc = 0;                                                                               \
            for (int64_t l = 0; l < K; ++l) acc += P[i * K + l] * Q[l * N + j];                           \
            T* c = (layout == 'C') ? &C[i + j * ldc] : &C[i * ldc + j];                                   \
            double prev = (beta == 0) ? 0.0 : beta * (double)(*c);                                        \
            *c = (T)(alpha * acc + prev);                                                                 \
        }                                                                                                 \
}
DEF_GEMM_DEF(f32, float)
DEF_GEMM_DEF(f64, double)

/* op(X) (R x C after op) from a `layout` buffer with leading dimension ld -> row-major double. */
#define DEF_DENSIFY(SUF, T)                                                                             \
static double* densify_##SUF(char layout, char op, int64_t R, int64_t C, const T* X, int64_t ld) {       \
    double* out = (double*) malloc(sizeof(double) * (size_t)(R * C > 0 ? R * C : 1));                     \
    for (int64_t i = 0; i < R; ++i) for (int64_t j = 0; j < C; ++j) {                                     \
        int64_t r = (op == 'N') ? i : j, c = (op == 'N') ? j : i;                                         \
        out[i * C + j] = (double)((layout == 'C') ? X[r + c * ld] : X[r * ld + c]);                       \
    }                                                                                                     \
    return out;                                                                                           \
}
DEF_DENSIFY(f32, float)
DEF_DENSIFY(f64, double)

/* op(S[ro:ro+.., co:co+..]) (R x C after op) of a dense-distribution operator, as T-valued doubles. */
static double* dense_op_block(int is_f32, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                              const uint32_t ctr[4], const uint32_t key[2], char op, int64_t R, int64_t C,
                              int64_t ro_s, int64_t co_s) {
    orc_state st; memcpy(st.c, ctr, 16); memcpy(st.k, key, 8);
    double* out = (double*) malloc(sizeof(double) * (size_t)(R * C > 0 ? R * C : 1));
    #pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < R; ++i) for (int64_t j = 0; j < C; ++j) {
        int64_t r = (op == 'N') ? i : j, c = (op == 'N') ? j : i;
        float v = dist_entry_raw(D_rows, D_cols, family, major_axis, &st, ro_s + r, co_s + c);
        double dv;
        if (family == 'U') dv = is_f32 ? (double)(v * (float) sqrt(3.0)) : (double) v * sqrt(3.0);
        else dv = (double) v;
        out[i * C + j] = dv;
    }
    return out;
}

#define DEF_SK_DENSE(SUF, T, ISF32)                                                                     \
int orc_sketch_dense_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, \
                           T alpha, int64_t D_rows, int64_t D_cols, char family, char major_axis,         \
                           const uint32_t ctr[4], const uint32_t key[2], int64_t ro_s, int64_t co_s,      \
                           const T* A, int64_t lda, T beta, T* B, int64_t ldb) {                          \
    if (side == 'L') {                                                                                    \
        int64_t d = p0, n = p1, m = p2;                                                                   \
        int64_t rs = (opS == 'N') ? d : m, cs = (opS == 'N') ? m : d;                                     \
        if (D_rows < rs + ro_s || D_cols < cs + co_s) return 1;                                           \
        double* P = dense_op_block(ISF32, D_rows, D_cols, family, major_axis, ctr, key, opS, d, m, ro_s, co_s); \
        double* Q = densify_##SUF(layout, opA, m, n, A, lda);                                             \
        gemm_def_##SUF(layout, d, n, m, (double) alpha, P, Q, (double) beta, B, ldb);                     \
        free(P); free(Q);                                                                                 \
    } else {                                                                                              \
        int64_t m = p0, d = p1, n = p2;                                                                   \
        int64_t rs = (opS == 'N') ? n : d, cs = (opS == 'N') ? d : n;                                     \
        if (D_rows < rs + ro_s || D_cols < cs + co_s) return 1;                                           \
        double* P = densify_##SUF(layout, opA, m, n, A, lda);                                             \
        double* Q = dense_op_block(ISF32, D_rows, D_cols, family, major_axis, ctr, key, opS, n, d, ro_s, co_s); \
        gemm_def_##SUF(layout, m, d, n, (double) alpha, P, Q, (double) beta, B, ldb);                     \
        free(P); free(Q);                                                                                 \
    }                                                                                                     \
    return 0;                                                                                             \
}
DEF_SK_DENSE(f32, float, 1)
DEF_SK_DENSE(f64, double, 0)

/* op(S[ro:, co:]) block of a sparse operator given as COO triplets (any order; duplicates add). */
#define DEF_COO_BLOCK(SUF, T)                                                                           \
static double* coo_op_block_##SUF(int64_t nnz, const int64_t* rows, const int64_t* cols, const T* vals,  \
                                  char op, int64_t R, int64_t C, int64_t ro, int64_t co) {                \
    double* out = (double*) calloc((size_t)(R * C > 0 ? R * C : 1), sizeof(double));                      \
    int64_t rs = (op == 'N') ? R : C, cs = (op == 'N') ? C : R;                                           \
    for (int64_t e = 0; e < nnz; ++e) {                                                                   \
        int64_t r = rows[e] - ro, c = cols[e] - co;                                                       \
        if (r < 0 || r >= rs || c < 0 || c >= cs) continue;                                               \
        if (op == 'N') out[r * C + c] += (double) vals[e]; else out[c * C + r] += (double) vals[e];       \
    }                                                                                                     \
    return out;                                                                                           \
}
DEF_COO_BLOCK(f32, float)
DEF_COO_BLOCK(f64, double)

#define DEF_SK_SPARSEOP(SUF, T)                                                                         \
int orc_sketch_sparseop_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, \
                              T alpha, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis,  \
                              const uint32_t ctr[4], const uint32_t key[2], int64_t ro_s, int64_t co_s,   \
                              const T* A, int64_t lda, T beta, T* B, int64_t ldb) {                       \
    int64_t lng = S_rows > S_cols ? S_rows : S_cols, sht = S_rows < S_cols ? S_rows : S_cols;             \
    int64_t nnz = vec_nnz * (major_axis == 'S' ? lng : sht);                                              \
    int64_t* rows = (int64_t*) malloc(8 * (size_t) nnz); int64_t* cols = (int64_t*) malloc(8 * (size_t) nnz); \
    T* vals = (T*) malloc(sizeof(T) * (size_t) nnz);                                                      \
    int rc = orc_fill_sparse_##SUF##_i64(S_rows, S_cols, vec_nnz, major_axis, ctr, key, rows, cols, vals, 0); \
    if (rc == 0) {                                                                                        \
        if (side == 'L') {                                                                                \
            int64_t d = p0, n = p1, m = p2;                                                               \
            double* P = coo_op_block_##SUF(nnz, rows, cols, vals, opS, d, m, ro_s, co_s);                 \
            double* Q = densify_##SUF(layout, opA, m, n, A, lda);                                         \
            gemm_def_##SUF(layout, d, n, m, (double) alpha, P, Q, (double) beta, B, ldb);                 \
            free(P); free(Q);                                                                             \
        } else {                                                                                          \
            int64_t m = p0, d = p1, n = p2;                                                               \
            double* P = densify_##SUF(layout, opA, m, n, A, lda);                                         \
            double* Q = coo_op_block_##SUF(nnz, rows, cols, vals, opS, n, d, ro_s, co_s);                 \
            gemm_def_##SUF(layout, m, d, n, (double) alpha, P, Q, (double) beta, B, ldb);                 \
            free(P); free(Q);                                                                             \
        }                                                                                                 \
    }                                                                                                     \
    free(rows); free(cols); free(vals);                                                                   \
    return rc;                                                                                            \
}
DEF_SK_SPARSEOP(f32, float)
DEF_SK_SPARSEOP(f64, double)

/* sketch_sparse: dense operator x sparse data. fmt 'O' COO (idx0 rows, idx1 cols), 'R' CSR
 * (idx0 rowptr, idx1 colidxs), 'C' CSC (idx0 rowidxs, idx1 colptr). */
#define DEF_SK_SPARSEDATA(SUF, T, ISF32)                                                                \
int orc_sketch_sparsedata_##SUF(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, \
                                T alpha, int64_t D_rows, int64_t D_cols, char family, char major_axis,    \
                                const uint32_t ctr[4], const uint32_t key[2], int64_t ro_s, int64_t co_s, \
                                char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz, const T* vals,     \
                                const int64_t* idx0, const int64_t* idx1, int64_t ro_a, int64_t co_a,     \
                                T beta, T* B, int64_t ldb) {                                              \
    int64_t* rows = (int64_t*) malloc(8 * (size_t)(nnz > 0 ? nnz : 1));                                   \
    int64_t* cols = (int64_t*) malloc(8 * (size_t)(nnz > 0 ? nnz : 1));                                   \
    if (fmt == 'O') { memcpy(rows, idx0, 8 * (size_t) nnz); memcpy(cols, idx1, 8 * (size_t) nnz); }       \
    else if (fmt == 'R') { for (int64_t i = 0; i < A_rows; ++i) for (int64_t e = idx0[i]; e < idx0[i + 1]; ++e) { rows[e] = i; cols[e] = idx1[e]; } } \
    else { for (int64_t j = 0; j < A_cols; ++j) for (int64_t e = idx1[j]; e < idx1[j + 1]; ++e) { cols[e] = j; rows[e] = idx0[e]; } } \
    if (side == 'L') {                                                                                    \
        int64_t d = p0, n = p1, m = p2;                                                                   \
        double* P = dense_op_block(ISF32, D_rows, D_cols, family, major_axis, ctr, key, opS, d, m, ro_s, co_s); \
        double* Q = coo_op_block_##SUF(nnz, rows, cols, vals, opA, m, n, ro_a, co_a);                     \
        gemm_def_##SUF(layout, d, n, m, (double) alpha, P, Q, (double) beta, B, ldb);                     \
        free(P); free(Q);                                                                                 \
    } else {                                                                                              \
        int64_t m = p0, d = p1, n = p2;                                                                   \
        double* P = coo_op_block_##SUF(nnz, rows, cols, vals, opA, m, n, ro_a, co_a);                     \
        double* Q = dense_op_block(ISF32, D_rows, D_cols, family, major_axis, ctr, key, opS, n, d, ro_s, co_s); \
        gemm_def_##SUF(layout, m, d, n, (double) alpha, P, Q, (double) beta, B, ldb);                     \
        free(P); free(Q);                                                                                 \
    }                                                                                                     \
    free(rows); free(cols);                                                                               \
    return 0;                                                                                             \
}
DEF_SK_SPARSEDATA(f32, float, 1)
DEF_SK_SPARSEDATA(f64, double, 0)

/* ------------------------------------------------------------------------------------------
 * Index sampling utilities (util.hh:257-334).
 * sample_indices_iid_uniform (:311-334) / sample_indices_iid (:284-305): the loop increments the counter BEFORE the first
 * draw ((i + 1) % 4 == 1 at i = 0), so sample i uses lane i % 4 of uneg11(philox(ctr + 1 + i / 4)); the value is mapped to
 * [0, 1] as (x + 1) / 2 in TF precision (double for the uniform form), then index = trunc(n * u) (uniform: `(int64_t) dN *
 * u` parses as ((int64_t) dN) * u, a double product truncated on assignment) or lower_bound(cdf, u).
 * Returned state: ctr + ceil(k / 4) for k > 0, ctr for k = 0.
 * weights_to_cdf (:257-270): sequential running sum in T (order matters), negatives >= error_if_below clipped to 0,
 * then scaled by 1 / sum with blas::scal; rc 1 where the reference throws.
 * ---------------------------------------------------------------------------------------- */
static float orc_sample_lane(const uint32_t ctr[4], const uint32_t key[2], int64_t i) {
    uint32_t c[4] = { ctr[0], ctr[1], ctr[2], ctr[3] }, rv[4];
    orc_ctr_add(c, (uint64_t)(1 + i / 4));
    orc_philox4x32_10(c, key, rv);
    return orc_uneg11(rv[i % 4]);
}
static void orc_sample_next(const uint32_t ctr[4], int64_t k, uint32_t next_ctr[4]) {
    if (!next_ctr) return;
    for (int q = 0; q < 4; ++q) next_ctr[q] = ctr[q];
    if (k > 0) orc_ctr_add(next_ctr, (uint64_t)((k + 3) / 4));
}
int orc_sample_indices_iid_uniform(int64_t n, int64_t k, int64_t* samples, const uint32_t ctr[4], const uint32_t key[2],
                                   uint32_t next_ctr[4]) {
    const double dN = (double) n;
    for (int64_t i = 0; i < k; ++i) {
        const double u = ((double) orc_sample_lane(ctr, key, i) + 1.0) / 2.0;
        samples[i] = (int64_t)((double)(int64_t) dN * u);
    }
    orc_sample_next(ctr, k, next_ctr);
    return 0;
}
#define DEF_SAMPLE(SUF, T)                                                                              \
int orc_sample_indices_iid_##SUF(int64_t n, const T* cdf, int64_t k, int64_t* samples, const uint32_t ctr[4], \
                                 const uint32_t key[2], uint32_t next_ctr[4]) {                           \
    for (int64_t i = 0; i < k; ++i) {                                                                     \
        const T u = ((T) orc_sample_lane(ctr, key, i) + (T) 1.0) / (T) 2.0;                               \
        int64_t lo = 0, hi = n;               /* std::lower_bound: first index with cdf[idx] >= u */     \
        while (lo < hi) { const int64_t mid = lo + (hi - lo) / 2; if (cdf[mid] < u) lo = mid + 1; else hi = mid; } \
        samples[i] = lo;                                                                                  \
    }                                                                                                     \
    orc_sample_next(ctr, k, next_ctr);                                                                    \
    return 0;                                                                                             \
}                                                                                                         \
int orc_weights_to_cdf_##SUF(int64_t n, T* w, T error_if_below) {                                         \
    T sum = (T) 0.0;                                                                                      \
    for (int64_t i = 0; i < n; ++i) {                                                                     \
        T val = w[i];                                                                                     \
        if (!(val >= error_if_below)) return 1;                                                           \
        val = val > (T) 0.0 ? val : (T) 0.0;                                                              \
        sum += val;                                                                                       \
        w[i] = sum;                                                                                       \
    }                                                                                                     \
    if (!(sum >= ((T) sqrt((double) n)) * (sizeof(T) == 4 ? (T) 0x1p-23 : (T) 0x1p-52))) return 1;        \
    const T inv = ((T) 1.0) / sum;                                                                        \
    for (int64_t i = 0; i < n; ++i) w[i] = inv * w[i];                                                    \
    return 0;                                                                                             \
}
DEF_SAMPLE(f32, float)
DEF_SAMPLE(f64, double)
