// Shared-memory wavefronts per warp instruction for the access shapes the SASO tile kernel uses or could use.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_wavefronts lds_wavefronts.cu
// Run under: ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum ./lds_wavefronts
// (one warp; wavefronts / executed shared-load instructions = wavefronts per instruction)
// Measured on B200, LDS.128: all lanes one address 2.06, one address per half-warp 2.06, one per quarter-warp 2.06,
// 32 distinct 16-byte pieces 4.06 -- a 16-byte shared load never costs fewer than two wavefronts, so sharing one slot
// record between the two half-warps would not make the record loads cheaper.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int ITER = 4096;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }

template <int BYTES> __device__ __forceinline__ uint32_t lds(uint32_t addr) {
    uint32_t a, b, c, d;
    if (BYTES == 16) { asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory"); return a + b + c + d; }
    if (BYTES == 8)  { asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr) : "memory"); return a + b; }
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(a) : "r"(addr) : "memory"); return a;
}
// SHAPE: 0 all lanes one address; 1 each half-warp its own address (adjacent); 2 lanes of a half contiguous, both halves the
// same addresses; 3 lanes of a half contiguous, halves in different rows (4 KB apart); 4 all 32 lanes contiguous;
// 5 each quarter-warp its own address (adjacent)
template <int BYTES, int SHAPE>
__global__ void k(uint32_t* out) {
    __shared__ __align__(128) uint8_t buf[16384];
    for (int i = threadIdx.x; i < 4096; i += 32) reinterpret_cast<uint32_t*>(buf)[i] = i;
    __syncthreads();
    const int lane = threadIdx.x, half = lane >> 4, l16 = lane & 15;
    uint32_t off;
    if (SHAPE == 0) off = 0;
    else if (SHAPE == 1) off = half * BYTES;
    else if (SHAPE == 2) off = l16 * BYTES;
    else if (SHAPE == 3) off = l16 * BYTES + half * 4096 + half * 136;
    else if (SHAPE == 4) off = lane * BYTES;
    else off = (lane >> 3) * BYTES;
    const uint32_t base = smem_u32(buf) + off;
    uint32_t acc = 0;
#pragma unroll 16
    for (int i = 0; i < ITER; ++i) acc += lds<BYTES>(base + (uint32_t)((i & 7) * 1024));
    out[threadIdx.x] = acc;
}
int main() {
    uint32_t* out; cudaMalloc(&out, 128);
    k<16, 0><<<1, 32>>>(out); k<16, 1><<<1, 32>>>(out); k<16, 5><<<1, 32>>>(out); k<16, 4><<<1, 32>>>(out);
    // (the 8- and 4-byte instantiations are not launched: nvcc hoists their loads out of the loop and the counts mean nothing)
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
