/* rbx.h -- C ABI of the B200 (sm_100a) sketching kernels behind the RandBLAS C++ surface.
 *
 * Every entry point is what a RandBLAS maintainer would bind in place of the body of one reference
 * function (file:line under the RandBLAS tree is cited at each declaration; INTEGRATION.md shows the
 * header-side stub). Plain pointers and sizes only; no C++ or torch types cross this boundary.
 *
 * Conventions
 *   layout 'C' (ColMajor) / 'R' (RowMajor); op 'N' / 'T'; side 'L' / 'R';
 *   family 'G' (Gaussian) / 'U' (uniform, scaled by sqrt(3) in fill_dense like dense_skops.hh:511-513);
 *   major axis 'S' (Short) / 'L' (Long)  -- same characters as RandBLAS::MajorAxis;
 *   RNG state = uint32 ctr[4] (128-bit little-endian counter) + uint32 key[2]  (base.hh:161-232);
 *   `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *   Pointers named *_dev are device pointers; the rbx_host_* variants take host pointers and stage
 *   through device memory (copies are part of the call).
 *   Return value: RBX_OK or a negative RBX_ERR_* code; rbx_strerror() names it. No exceptions cross.
 *   There is no CPU fallback: without a CUDA device every compute entry point returns RBX_ERR_NO_DEVICE.
 */
#ifndef RBX_H
#define RBX_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define RBX_OK              0
#define RBX_ERR_ARG        -1   /* a precondition the reference checks with randblas_require failed */
#define RBX_ERR_CUDA       -2   /* a CUDA runtime call or kernel launch failed */
#define RBX_ERR_NO_DEVICE  -3   /* no sm_100 device visible */
#define RBX_ERR_ALLOC      -4   /* device workspace allocation failed */
#define RBX_ERR_UNSUPPORTED -5  /* valid in the reference, not implemented on the device yet */
#define RBX_ERR_COMM       -7   /* NCCL missing (libnccl.so.2 could not be loaded) or an NCCL call failed */
#define RBX_ERR_TOO_LARGE  -6   /* beyond the fixed capacity of an allocation-free entry point (rbx_sketch_vector_noalloc) */

const char* rbx_strerror(int code);
const char* rbx_version(void);
int rbx_device_count(void);
/* number of kernel launches issued by this library in this process so far (for bench accounting) */
int64_t rbx_launch_count(void);

/* ---- device memory for callers that do not link the CUDA runtime (the header layer in include/RandBLAS_b200.hh) ---- */
int rbx_malloc(void** dev_ptr, uint64_t bytes);
int rbx_free(void* dev_ptr);
int rbx_memcpy_h2d(void* dst_dev, const void* src_host, uint64_t bytes, void* stream);
int rbx_memcpy_d2h(void* dst_host, const void* src_dev, uint64_t bytes, void* stream);
int rbx_stream_synchronize(void* stream);
/* Kernel workspaces (operator codes, split partials, operator slabs) come from a private stream-ordered pool per device
   (never the default pool the application shares) that keeps up to 1/8 of the device's memory mapped between calls (only what calls really used; re-mapping the 12 GB of a full-size sketch_sparse costs more than the call)
   (RBX_WS_KEEP_MB overrides); this synchronises the device and returns all of it to the driver. */
int rbx_trim_workspace(void);
/* bytes of the current device's private workspace pool: in use now, the most in use at any time since the last reset (the
   high-water mark: what a call really needed), and mapped now. reset_high != 0 restarts the high-water mark after reading it.
   No counterpart in the reference (its workspaces are std::vector / new[]); the tests use it to hold the implicit dense
   sketches to "no allocation proportional to d * m". */
int rbx_workspace_stats(uint64_t* used_now, uint64_t* used_high, uint64_t* reserved_now, int reset_high);

/* ---- RNG state algebra (host-side, exact) ------------------------------------------------------ */
/* ctr += n with carry: Random123 array.h incr(); pinned by test/test_basic_rng/test_r123.cc:679-766 */
void rbx_ctr_add(uint32_t ctr[4], uint64_t n);
/* dense::compute_next_state, dense_skops.hh:172-191 */
void rbx_dense_next_state(int64_t n_rows, int64_t n_cols, char major_axis, const uint32_t ctr[4], uint32_t next_ctr[4]);
/* sparse::compute_next_state, sparse_skops.hh:115-126 (reproduced as written there) */
void rbx_sparse_next_state(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], uint32_t next_ctr[4]);
/* dist_to_layout / major_axis_length, dense_skops.hh:297-316 */
char rbx_dist_to_layout(int64_t n_rows, int64_t n_cols, char major_axis);
int64_t rbx_major_axis_length(int64_t n_rows, int64_t n_cols, char major_axis);

/* ---- raw generator (testing the device Philox / transforms against KATs) ------------------------ */
/* out_dev[4*i..4*i+3] = Philox4x32-10(ctr + i, key), i < n.  r123_kat_vectors.txt:19-21 */
int rbx_philox4x32_dev(const uint32_t ctr[4], const uint32_t key[2], int64_t n, uint32_t* out_dev, void* stream);
/* r123ext::boxmul / uneg11 ::generate on (ctr + i, key): random_gen.hh:136-144,163-172 */
int rbx_generate4_dev(char family, const uint32_t ctr[4], const uint32_t key[2], int64_t n, float* out_dev, void* stream);
/* the float transforms on caller-chosen raw words (exhaustive sweeps against the host libm):
 * (s,c)[i] = boxmuller(u0[i], u1[i]) */
int rbx_boxmuller_raw_dev(int64_t n, const uint32_t* u0_dev, const uint32_t* u1_dev, float* s_dev, float* c_dev, void* stream);

/* ---- dense operators ----------------------------------------------------------------------------- */
/* dense::fill_dense_submat_impl<T,RNG,OP>, dense_skops.hh:96-170 */
int rbx_fill_dense_submat_f32(char family, int64_t n_cols, float* smat_dev, int64_t n_srows, int64_t n_scols,
                              int64_t ptr, const uint32_t ctr[4], const uint32_t key[2], int64_t lda,
                              uint32_t next_ctr[4], void* stream);
int rbx_fill_dense_submat_f64(char family, int64_t n_cols, double* smat_dev, int64_t n_srows, int64_t n_scols,
                              int64_t ptr, const uint32_t ctr[4], const uint32_t key[2], int64_t lda,
                              uint32_t next_ctr[4], void* stream);
/* fill_dense(layout, D, n_rows, n_cols, ro_s, co_s, buff, seed), dense_skops.hh:486-532 */
int rbx_fill_dense_f32(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                       int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, float* buff_dev,
                       const uint32_t ctr[4], const uint32_t key[2], uint32_t next_ctr[4], void* stream);
int rbx_fill_dense_f64(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                       int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, double* buff_dev,
                       const uint32_t ctr[4], const uint32_t key[2], uint32_t next_ctr[4], void* stream);
int rbx_host_fill_dense_f32(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                            int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, float* buff_host,
                            const uint32_t ctr[4], const uint32_t key[2], uint32_t next_ctr[4]);
int rbx_host_fill_dense_f64(char layout, int64_t D_rows, int64_t D_cols, char family, char major_axis,
                            int64_t n_rows, int64_t n_cols, int64_t ro_s, int64_t co_s, double* buff_host,
                            const uint32_t ctr[4], const uint32_t key[2], uint32_t next_ctr[4]);

/* ---- sparse operators ---------------------------------------------------------------------------- */
/* sparse::repeated_fisher_yates, sparse_skops.hh:53-113 (idxs_minor_dev / vals_dev may be NULL) */
int rbx_repeated_fisher_yates_f32_i32(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz, int64_t dim_major,
                                      int64_t dim_minor, int32_t* idxs_major_dev, int32_t* idxs_minor_dev, float* vals_dev, void* stream);
int rbx_repeated_fisher_yates_f32_i64(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz, int64_t dim_major,
                                      int64_t dim_minor, int64_t* idxs_major_dev, int64_t* idxs_minor_dev, float* vals_dev, void* stream);
int rbx_repeated_fisher_yates_f64_i32(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz, int64_t dim_major,
                                      int64_t dim_minor, int32_t* idxs_major_dev, int32_t* idxs_minor_dev, double* vals_dev, void* stream);
int rbx_repeated_fisher_yates_f64_i64(const uint32_t ctr[4], const uint32_t key[2], int64_t vec_nnz, int64_t dim_major,
                                      int64_t dim_minor, int64_t* idxs_major_dev, int64_t* idxs_minor_dev, double* vals_dev, void* stream);
/* fill_sparse(SparseSkOp&), sparse_skops.hh:389-413: COO triplets rows/cols/vals of length
 * vec_nnz * (long dim) for Short-axis (SASO) or vec_nnz * (short dim) for Long-axis (LASO) */
int rbx_fill_sparse_f32_i32(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int32_t* rows_dev, int32_t* cols_dev, float* vals_dev, uint32_t next_ctr[4], void* stream);
int rbx_fill_sparse_f32_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int64_t* rows_dev, int64_t* cols_dev, float* vals_dev, uint32_t next_ctr[4], void* stream);
int rbx_fill_sparse_f64_i32(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int32_t* rows_dev, int32_t* cols_dev, double* vals_dev, uint32_t next_ctr[4], void* stream);
int rbx_fill_sparse_f64_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int64_t* rows_dev, int64_t* cols_dev, double* vals_dev, uint32_t next_ctr[4], void* stream);
int rbx_host_fill_sparse_f64_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t* rows_host, int64_t* cols_host, double* vals_host, uint32_t next_ctr[4]);
int rbx_host_fill_sparse_f32_i64(int64_t n_rows, int64_t n_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t* rows_host, int64_t* cols_host, float* vals_host, uint32_t next_ctr[4]);

/* ---- sketching with a dense operator: dense::lskge3 / dense::rskge3, skge.hh:173-215 / 320-364 ---- */
/* side 'L': B(d x n) = alpha * op(submat(S)) * op(A) + beta * B, dims (p0,p1,p2) = (d, n, m)
 * side 'R': B(m x d) = alpha * op(A) * op(submat(S)) + beta * B, dims (p0,p1,p2) = (m, d, n)
 * submat(S) starts at (ro_s, co_s) of the D_rows x D_cols operator defined by (family, major_axis, ctr, key).
 * S_buff_dev == NULL: the operator is generated inside the kernel and never written to memory.
 * S_buff_dev != NULL: explicit operator (DenseSkOp::buff) stored in S_layout with the natural leading
 * dimension (D_cols for 'R', D_rows for 'C'), as after fill_dense(DenseSkOp&), dense_skops.hh:583-592. */
int rbx_sketch_dense_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                         int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                         const float* S_buff_dev, char S_layout, int64_t ro_s, int64_t co_s,
                         const float* A_dev, int64_t lda, float beta, float* B_dev, int64_t ldb, void* stream);
int rbx_sketch_dense_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                         int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                         const double* S_buff_dev, char S_layout, int64_t ro_s, int64_t co_s,
                         const double* A_dev, int64_t lda, double beta, double* B_dev, int64_t ldb, void* stream);
int rbx_host_sketch_dense_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                              int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                              int64_t ro_s, int64_t co_s, const float* A_host, int64_t lda, float beta, float* B_host, int64_t ldb);
int rbx_host_sketch_dense_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                              int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                              int64_t ro_s, int64_t co_s, const double* A_host, int64_t lda, double beta, double* B_host, int64_t ldb);

/* Allocation-free sketch_vector: y (d') = alpha * op(submat(S)) * x + beta * y for an implicit dense operator, dtype 'f'
 * (float) or 'd' (double) for x / y / alpha / beta (alpha and beta are passed as double and narrowed); (d, m) are the dims of
 * submat(S) BEFORE op, as in sketch_vector (skve.hh:152-176). Same kernels and results as rbx_sketch_dense_* with n = 1
 * (gemv_rows / gemv_cols + gemv_reduce), but the per-chunk partial sums live in a file-scope `static __device__` array of
 * 4 MiB instead of a stream-ordered allocation: the call issues no cudaMalloc / cudaMallocAsync / cudaFree (usable under
 * CUDA graph capture and in loops that may not allocate). Limit: ceil(K / c) * M * sizeof(T) <= 4 MiB, where M / K are the
 * output / contraction lengths and c = 1024 (generator rows along the output) or 64 (along the contraction); beyond it the
 * call returns RBX_ERR_TOO_LARGE and prints the sizes. All calls on one GPU share the array: calls that overlap on the GPU
 * (different streams) race; calls ordered on one stream are safe. Nothing in the array has to start at zero.
 * NO COUNTERPART IN THE REFERENCE. */
int rbx_sketch_vector_noalloc(char dtype, char opS, int64_t d, int64_t m, double alpha,
                              int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                              int64_t ro_s, int64_t co_s, const void* x_dev, int64_t incx, double beta, void* y_dev, int64_t incy,
                              void* stream);

/* ---- sketching with a sparse operator: sparse::lskges / sparse::rskges, skge.hh:485-510 / 616-641 -- */
/* Implicit operator: the SASO/LASO defined by (S_rows, S_cols, vec_nnz, major_axis, ctr, key) is regenerated on
 * the device (fill_sparse semantics, sparse_skops.hh:389-413); nothing proportional to nnz is returned. */
int rbx_sketch_sparseop_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                            int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int64_t ro_s, int64_t co_s, const float* A_dev, int64_t lda, float beta, float* B_dev, int64_t ldb, void* stream);
int rbx_sketch_sparseop_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                            int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                            int64_t ro_s, int64_t co_s, const double* A_dev, int64_t lda, double beta, double* B_dev, int64_t ldb, void* stream);
int rbx_host_sketch_sparseop_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                                 int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t ro_s, int64_t co_s, const float* A_host, int64_t lda, float beta, float* B_host, int64_t ldb);
int rbx_host_sketch_sparseop_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                                 int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                 int64_t ro_s, int64_t co_s, const double* A_host, int64_t lda, double beta, double* B_host, int64_t ldb);
/* Explicit COO operator / COO sparse matrix times dense: left_spmm / right_spmm with a COOMatrix,
 * spmm_dispatch.hh:48-200 + coo_spmm_impl.hh:79-162 (int64 indices, index_base Zero). */
int rbx_spmm_coo_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                     int64_t S_rows, int64_t S_cols, int64_t nnz, const int64_t* rows_dev, const int64_t* cols_dev, const float* vals_dev,
                     int64_t ro_s, int64_t co_s, const float* A_dev, int64_t lda, float beta, float* B_dev, int64_t ldb, void* stream);
int rbx_spmm_coo_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                     int64_t S_rows, int64_t S_cols, int64_t nnz, const int64_t* rows_dev, const int64_t* cols_dev, const double* vals_dev,
                     int64_t ro_s, int64_t co_s, const double* A_dev, int64_t lda, double beta, double* B_dev, int64_t ldb, void* stream);

/* CSR ('R': ptr = rowptr, idx = colidxs) or CSC ('C': ptr = colptr, idx = rowidxs) operand times dense: left_spmm / right_spmm
 * with a CSRMatrix / CSCMatrix, spmm_dispatch.hh:48-200 + csr_spmm_impl.hh:45-156 / csc_spmm_impl.hh:43-213. The operand is
 * taken whole (dimensions equal to the product's, spmm_dispatch.hh:97-104). An operand compressed along the product's output
 * axis (CSR not transposed, CSC transposed) is used in place -- no regrouping, no copy; the other two cases regroup as
 * rbx_spmm_coo does. int64 indices, index_base Zero. */
int rbx_spmm_csx_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                     int64_t S_rows, int64_t S_cols, int64_t nnz, char fmt, const int64_t* ptr_dev, const int64_t* idx_dev, const float* vals_dev,
                     const float* A_dev, int64_t lda, float beta, float* B_dev, int64_t ldb, void* stream);
int rbx_spmm_csx_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                     int64_t S_rows, int64_t S_cols, int64_t nnz, char fmt, const int64_t* ptr_dev, const int64_t* idx_dev, const double* vals_dev,
                     const double* A_dev, int64_t lda, double beta, double* B_dev, int64_t ldb, void* stream);

/* ---- index sampling utilities, util.hh:257-334 (same Philox + uneg11 stream as the reference, bit for bit) ----
 * rbx_sample_indices_iid_uniform: util::sample_indices_iid_uniform (util.hh:311-334), k samples of {0..n-1};
 * rbx_sample_indices_iid_*:       util::sample_indices_iid (util.hh:284-305), lower_bound in a device-resident cdf;
 * rbx_weights_to_cdf_*:           util::weights_to_cdf (util.hh:257-270), in place; RBX_ERR_ARG where the reference throws
 *                                 (a weight below error_if_below, or a total below sqrt(n) * eps); synchronises the stream.
 * next_ctr (may be NULL) receives the returned state's counter: ctr + ceil(k / 4) for k > 0. */
int rbx_sample_indices_iid_uniform(int64_t n, int64_t k, int64_t* samples_dev, const uint32_t ctr[4], const uint32_t key[2],
                                   uint32_t next_ctr[4], void* stream);
int rbx_sample_indices_iid_f32(int64_t n, const float* cdf_dev, int64_t k, int64_t* samples_dev, const uint32_t ctr[4],
                               const uint32_t key[2], uint32_t next_ctr[4], void* stream);
int rbx_sample_indices_iid_f64(int64_t n, const double* cdf_dev, int64_t k, int64_t* samples_dev, const uint32_t ctr[4],
                               const uint32_t key[2], uint32_t next_ctr[4], void* stream);
int rbx_weights_to_cdf_f32(int64_t n, float* w_dev, float error_if_below, void* stream);
int rbx_weights_to_cdf_f64(int64_t n, double* w_dev, double error_if_below, void* stream);

/* ---- sparse format conversion on the device: coo_to_csc / coo_to_csr (conversions.hh:43-62, 84-103, which sort the COO
 * triplets on the host with sort_coo_data, coo_matrix.hh:267-325) and CSR <-> CSC. by 'C': compressed columns
 * (ptr_out = colptr, n_cols + 1 entries; idx_out = rowidxs); by 'R': compressed rows (ptr_out = rowptr, n_rows + 1;
 * idx_out = colidxs). Inner indices come out increasing inside every segment (duplicates ordered by value bits), so the
 * result is canonical whatever the input order. fmt_in / idx0 / idx1 as for rbx_sketch_sparsedata. Outputs are device
 * arrays of the caller (nnz entries each, n_outer + 1 for ptr_out). */
int rbx_sparse_regroup_f32(char by, char fmt_in, int64_t n_rows, int64_t n_cols, int64_t nnz, const float* vals_dev,
                           const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t* ptr_out_dev, int64_t* idx_out_dev,
                           float* vals_out_dev, void* stream);
int rbx_sparse_regroup_f64(char by, char fmt_in, int64_t n_rows, int64_t n_cols, int64_t nnz, const double* vals_dev,
                           const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t* ptr_out_dev, int64_t* idx_out_dev,
                           double* vals_out_dev, void* stream);

/* util::require_symmetric (util.hh:165-188) for a device-resident n x n matrix: RBX_ERR_ARG when some pair violates
 * |A_ij - A_ji| <= (|A_ij| + |A_ji| + 1) * tol (the reference throws "Symmetry check failed"); tol < 0 skips the check.
 * One kernel over the upper triangle and a 4-byte read-back; synchronises the stream. (sksy.hh:165-184 calls it before sketching) */
int rbx_require_symmetric_f32(char layout, const float* A_dev, int64_t n, int64_t lda, float tol, void* stream);
int rbx_require_symmetric_f64(char layout, const double* A_dev, int64_t n, int64_t lda, double tol, void* stream);

/* outer index of every nonzero of a CSR (rows) / CSC (columns) matrix from its pointer array: the device-side step of viewing
 * compressed data as COO triplets (left_spmm / right_spmm on CSR / CSC, spmm_dispatch.hh:48-200). outer_idx_dev has nnz =
 * ptr[n_outer] entries. */
int rbx_expand_ptr(int64_t n_outer, const int64_t* ptr_dev, int64_t* outer_idx_dev, void* stream);

/* ---- sketch_sparse: dense operator x sparse data, sparse_data::lsksp3 / rsksp3, sksp.hh:147-192 / 302-345 */
/* fmt 'O' COO (idx0 = rows, idx1 = cols), 'R' CSR (idx0 = rowptr, idx1 = colidxs), 'C' CSC (idx0 = rowidxs,
 * idx1 = colptr); int64 indices, index_base Zero. CSR/CSC require ro_a == co_a == 0 (spmm_dispatch.hh:99-104). */
int rbx_sketch_sparsedata_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                              int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                              int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                              const float* vals_dev, const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t ro_a, int64_t co_a,
                              float beta, float* B_dev, int64_t ldb, void* stream);
int rbx_sketch_sparsedata_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                              int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                              int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                              const double* vals_dev, const int64_t* idx0_dev, const int64_t* idx1_dev, int64_t ro_a, int64_t co_a,
                              double beta, double* B_dev, int64_t ldb, void* stream);

/* the same with every array in host memory (the arrays a caller of RandBLAS::sketch_sparse owns, sksp.hh:464-616): the sparse
 * matrix and B are staged through device memory around the call; synchronous */
int rbx_host_sketch_sparsedata_f32(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, float alpha,
                                   int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                   int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                                   const float* vals_host, const int64_t* idx0_host, const int64_t* idx1_host, int64_t ro_a, int64_t co_a,
                                   float beta, float* B_host, int64_t ldb);
int rbx_host_sketch_sparsedata_f64(char side, char layout, char opS, char opA, int64_t p0, int64_t p1, int64_t p2, double alpha,
                                   int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                   int64_t ro_s, int64_t co_s, char fmt, int64_t A_rows, int64_t A_cols, int64_t nnz,
                                   const double* vals_host, const int64_t* idx0_host, const int64_t* idx1_host, int64_t ro_a, int64_t co_a,
                                   double beta, double* B_host, int64_t ldb);

/* ---- multi-GPU (reference contract: rtd/source/tutorial/updates.rst:59-80, skge.hh:182-189) ----------------------
 * One process per GPU. rbx_shard_range: rank's [start, start+len) of `total`, in units of `align`.
 * Column-sharded sketches need no entry point: every rank calls rbx_sketch_* on its own column slab with the same operator.
 * Row-sharded left sketches: rbx_shard_sketch_*: rank r holds rows [start_r, start_r + len_r) of op(A)
 * (rbx_shard_range(m_total, world, rank, align)); it sketches them with the operator's column offset co_s + start_r (beta is
 * applied on rank 0 only, so beta * B enters the sum once), then ONE NCCL collective on `stream` sums the d x n partials:
 * ncclReduceScatter into B_piece_dev (d * n / world elements: rank r receives elements [r, r+1) * d * n / world of the
 * flattened B; d * n must divide by world) or, with B_piece_dev == NULL, ncclAllReduce in place. B must be dense
 * (ldb == d for 'C', n for 'R'). comm == NULL or world == 1: no collective (B_piece_dev, if given, receives a copy).
 * The communicator: rank 0 calls rbx_comm_unique_id and hands the 128 bytes to the other ranks by whatever channel the
 * application has (MPI, torch.distributed, a file); every rank then calls rbx_comm_init with its device current.
 * NCCL is loaded at run time (libnccl.so.2, or RBX_NCCL_LIBRARY); without it these calls return RBX_ERR_COMM. */
void rbx_shard_range(int64_t total, int world, int rank, int64_t align, int64_t* start, int64_t* len);
int rbx_comm_unique_id(uint8_t id[128]);
int rbx_comm_init(void** comm, int world, int rank, const uint8_t id[128]);
int rbx_comm_destroy(void* comm);
int rbx_comm_info(void* comm, int* world, int* rank, int* nccl_version);
int rbx_comm_reduce_scatter_f32(void* comm, float* B_dev, int64_t count, float* piece_dev, void* stream);
int rbx_comm_reduce_scatter_f64(void* comm, double* B_dev, int64_t count, double* piece_dev, void* stream);
int rbx_comm_all_reduce_f32(void* comm, float* B_dev, int64_t count, void* stream);
int rbx_comm_all_reduce_f64(void* comm, double* B_dev, int64_t count, void* stream);
/* dense operator (lskge3 per rank, skge.hh:173-215) */
int rbx_shard_sketch_dense_f32(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, float alpha,
                               int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                               int64_t ro_s, int64_t co_s, const float* A_local_dev, int64_t lda, float beta,
                               float* B_dev, int64_t ldb, float* B_piece_dev, int64_t align, void* stream);
int rbx_shard_sketch_dense_f64(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, double alpha,
                               int64_t D_rows, int64_t D_cols, char family, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                               int64_t ro_s, int64_t co_s, const double* A_local_dev, int64_t lda, double beta,
                               double* B_dev, int64_t ldb, double* B_piece_dev, int64_t align, void* stream);
/* sparse operator (lskges per rank, skge.hh:485-510): the columns [start_r, start_r + len_r) of a wide SASO are exactly the
 * Fisher-Yates vectors of those indices, so a rank regenerates only its own part of S */
int rbx_shard_sketch_sparseop_f32(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, float alpha,
                                  int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                  int64_t ro_s, int64_t co_s, const float* A_local_dev, int64_t lda, float beta,
                                  float* B_dev, int64_t ldb, float* B_piece_dev, int64_t align, void* stream);
int rbx_shard_sketch_sparseop_f64(void* comm, char layout, char opA, int64_t d, int64_t n, int64_t m_total, double alpha,
                                  int64_t S_rows, int64_t S_cols, int64_t vec_nnz, char major_axis, const uint32_t ctr[4], const uint32_t key[2],
                                  int64_t ro_s, int64_t co_s, const double* A_local_dev, int64_t lda, double beta,
                                  double* B_dev, int64_t ldb, double* B_piece_dev, int64_t align, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RBX_H */
