// K2 tensor-core paths (placeholder dispatch until the tcgen05 / DMMA kernels land in this file).
#include "rbx_kernels.cuh"
namespace rbx {
int launch_sketch_dense_tc_f32(const GemmOperands<float>&, cudaStream_t) { return RBX_ERR_UNSUPPORTED; }
int launch_sketch_dense_tc_f64(const GemmOperands<double>&, cudaStream_t) { return RBX_ERR_UNSUPPORTED; }
}
