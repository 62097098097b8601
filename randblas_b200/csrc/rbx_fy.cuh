// One minor-axis vector of sparse::repeated_fisher_yates (sparse_skops.hh:71-103), kept lazily: the work array is
// the identity except at the (at most vec_nnz) positions remembered in pos[]/val[].
#pragma once
#include "rbx_rng.cuh"

namespace rbx {

// emit(j, major_index, negative) for j = 0 .. vec_nnz-1 of vector i. Counter: seed + i*vec_nnz + j (:74-78, :91).
template <int KMAX, typename Emit>
__device__ __forceinline__ void fy_vector(const Ctr128& ctr, const Key64& key, int64_t i, int vec_nnz, int64_t dim_major, Emit emit) {
    int64_t pos[KMAX], val[KMAX];
    int nov = 0;
    const bool small = dim_major <= 0xffffffffll;
    for (int j = 0; j < vec_nnz; ++j) {
        const uint4 rv = philox4x32_10(ctr_add(ctr, (uint64_t)(i * vec_nnz + j)), key);
        const int64_t span = dim_major - j;
        // ell = j + rv[0] % (dim_major - j): rv[0] is promoted to int64 (sparse_skops.hh:80)
        const int64_t ell = j + (small ? (int64_t)(rv.x % (uint32_t) span) : (int64_t)((uint64_t) rv.x % (uint64_t) span));
        int64_t at_ell = ell, at_j = j;
        int slot = -1;
        for (int t = 0; t < nov; ++t) {
            if (pos[t] == ell) { at_ell = val[t]; slot = t; }
            if (pos[t] == j) at_j = val[t];
        }
        if (ell != j) {                       // work[ell] <- work[j]; work[j] is never read again
            if (slot < 0) { slot = nov++; pos[slot] = ell; }
            val[slot] = at_j;
        }
        emit(j, at_ell, (rv.y & 1u) != 0);    // vals = (rv[1] % 2 == 0) ? +1 : -1 (:87)
    }
}

}  // namespace rbx
