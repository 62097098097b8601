// Internal launcher prototypes shared between the kernel translation units and the C ABI (rbx_capi.cu).
#pragma once
#include "rbx_common.cuh"

namespace rbx {

void count_launch();

// ---- K1 -------------------------------------------------------------------------------------------
template <typename T>
int launch_fill_window(char family, int64_t n_cols_parent, T* out, int64_t n_srows, int64_t n_scols, int64_t r0, int64_t c0,
                       const Ctr128& ctr, const Key64& key, int64_t rs, int64_t cs, bool do_scale, cudaStream_t stream);
int launch_philox_stream(const Ctr128& ctr, const Key64& key, int64_t n, uint32_t* out, cudaStream_t stream);
int launch_generate4_stream(char family, const Ctr128& ctr, const Key64& key, int64_t n, float* out, cudaStream_t stream);
int launch_boxmuller_raw(int64_t n, const uint32_t* u0, const uint32_t* u1, float* s, float* c, cudaStream_t stream);

// ---- K3 -------------------------------------------------------------------------------------------
template <typename T, typename I>
int launch_fisher_yates(const Ctr128& ctr, const Key64& key, int64_t vec_nnz, int64_t dim_major, int64_t dim_minor,
                        I* idxs_major, I* idxs_minor, T* vals, cudaStream_t stream);

// ---- index sampling utilities (rbx_sample.cu; util.hh:257-334) ------------------------------------------
int launch_sample_uniform(int64_t n, int64_t k, int64_t* samples, const Ctr128& ctr, const Key64& key, cudaStream_t stream);
template <typename T>
int launch_sample_cdf(int64_t n, const T* cdf, int64_t k, int64_t* samples, const Ctr128& ctr, const Key64& key, cudaStream_t stream);
template <typename T> int launch_weights_to_cdf(int64_t n, T* w, T error_if_below, cudaStream_t stream);

// ---- K2 -------------------------------------------------------------------------------------------
// C[i, j] = alpha * sum_l Sop[i, l] * X[l, j] + beta * C[i, j]
template <typename T>
struct GemmOperands {
    T* C; int64_t sc_i, sc_j;              // M x N output, element strides
    const T* X; int64_t sx_l, sx_j;        // K x N data, element strides
    int64_t M, N, K;
    T alpha, beta;
    // implicit operator: Sop[i, l] = parent(pr0 + (l_on_cols ? i : l), pc0 + (l_on_cols ? l : i))
    DenseOpWindow w; int l_on_cols;
    // explicit operator (DenseSkOp::buff): Sop[i, l] = Sx[i * ss_i + l * ss_l]; nullptr => implicit
    const T* Sx; int64_t ss_i, ss_l;
    int64_t nbx;                           // column tiles (set by the launcher)
};
template <typename T> int launch_sketch_dense_simt(const GemmOperands<T>& g, cudaStream_t stream);
// one right-hand side (sketch_vector): generator + dot product; RBX_ERR_UNSUPPORTED unless N == 1 and the operator is implicit
template <typename T> int launch_sketch_dense_gemv(const GemmOperands<T>& g, cudaStream_t stream);
// the same kernels with their partial sums in a static __device__ array (no allocation); RBX_ERR_TOO_LARGE beyond its size
template <typename T> int launch_sketch_dense_gemv_static(const GemmOperands<T>& g, cudaStream_t stream);
// tensor-core paths; return RBX_ERR_UNSUPPORTED when the shape/layout is outside their envelope
int launch_sketch_dense_tc_f32(const GemmOperands<float>& g, cudaStream_t stream);
int launch_sketch_dense_tc_f64(const GemmOperands<double>& g, cudaStream_t stream);

// ---- K4 / K5 --------------------------------------------------------------------------------------
// C = beta * C over an M x N strided view (beta == 0 writes zeros without reading: util.hh:51-59)
template <typename T> int launch_scale_view(T* C, int64_t M, int64_t N, int64_t sc_i, int64_t sc_j, T beta, cudaStream_t stream);

// util::require_symmetric on the device (util.hh:165-188): raises *violated_dev when some |A_ij - A_ji| exceeds its tolerance
template <typename T>
int launch_require_symmetric(const T* A, int64_t n, int64_t s_row, int64_t s_col, T tol, int* violated_dev, cudaStream_t stream);

// K5 (rbx_spdata.cu): output-stationary gather. Column j of C gets alpha * sum over the entries (l, v) of segment j of
// v * Sd[(l - l0) * Mp + i]; `first` applies beta (C not read when beta == 0), later slabs accumulate.
template <typename T>
int launch_spdata_gather(int64_t N, const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t l0, int64_t kk, int64_t M, int64_t Mp,
                         const float* Sd, bool uniform, bool first, T alpha, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream);
// the same product for `rows` <= dc output rows whose operator slab Sd[l * dc + i] (dc in {8, 16, 32}) is small enough to stay
// in L2 while the nonzeros stream through (d-blocked passes of sketch_sparse); applies alpha / beta at its single store
template <typename T>
int launch_spdata_gather_narrow(int64_t N, const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t K, int rows, int dc, const float* Sd,
                                bool uniform, T alpha, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream);
// regroup per-nonzero (p, q, v) into column segments of op(A) (sorted by l inside a segment). The three outputs are
// stream-ordered allocations the caller releases with cudaFreeAsync.
template <typename T>
int spd_group_by_column(int64_t nnz, const int64_t* prow, const int64_t* pcol, const T* vals, bool trans, int64_t ro, int64_t co,
                        int64_t K, int64_t N, int64_t** ptr_out, int64_t** lidx_out, T** vals_out, cudaStream_t stream,
                        int lc_shift = 62, int64_t nlc = 1);   // nlc > 1: sub-segments per block of 2^lc_shift contraction indices

// Deterministic general sparse-operator product (rbx_spmm_gather.cu): C = alpha * Sop * X + beta * C for COO triplets,
// (i, l) = trans ? (c - co, r - ro) : (r - ro, c - co); entries are regrouped by output row on the device and every output
// element is summed by one warp in a fixed order -- no atomics. beta is applied here (C is not read when beta == 0).
template <typename T>
int launch_coo_gather(int64_t nnz, const int64_t* rows, const int64_t* cols, const T* vals, bool trans, int64_t ro, int64_t co,
                      int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                      T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream);

// the same product for an operand already compressed along the output axis (its ptr / idx / vals arrays are the segments)
template <typename T>
int launch_csx_gather(const int64_t* ptr, const int64_t* lidx, const T* vals, int64_t nnz, int64_t M, int64_t N, int64_t K, T alpha,
                      const T* X, int64_t sx_l, int64_t sx_j, T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream);

// expand a CSR rowptr / CSC colptr (n_outer + 1 entries) into one outer index per nonzero
int launch_expand_ptr(int64_t n_outer, const int64_t* ptr, int64_t* outer_idx, cudaStream_t stream);

// fast SASO path (left sketch, wide Short-axis operator, contiguous rows of X); RBX_ERR_UNSUPPORTED otherwise
template <typename T>
int launch_saso_fast(const Ctr128& ctr, const Key64& key, int64_t S_rows, int64_t S_cols, int64_t vec_nnz, int64_t ro, int64_t co,
                     int64_t M, int64_t N, int64_t K, T alpha, const T* X, int64_t sx_l, int64_t sx_j,
                     T beta, T* C, int64_t sc_i, int64_t sc_j, cudaStream_t stream);

}  // namespace rbx
