// K4 streaming SASO kernel (placeholder dispatch until the kernel lands in this file).
#include "rbx_kernels.cuh"
namespace rbx {
template <typename T>
int launch_saso_fast(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, T,
                     const T*, int64_t, int64_t, T, T*, int64_t, int64_t, cudaStream_t) { return RBX_ERR_UNSUPPORTED; }
template int launch_saso_fast<float>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, float,
                     const float*, int64_t, int64_t, float, float*, int64_t, int64_t, cudaStream_t);
template int launch_saso_fast<double>(const Ctr128&, const Key64&, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t, double,
                     const double*, int64_t, int64_t, double, double*, int64_t, int64_t, cudaStream_t);
}
