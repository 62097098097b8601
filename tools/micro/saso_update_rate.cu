// Update-rate ceilings of the candidate SASO inner loops, per SM, measured.
// One CTA per SM (512 threads, 148 CTAs), all operands on chip, no global traffic in the timed loops:
//   G32  register-resident B rows (static index), operands GATHERED from a shared-memory A chunk, 32-bit codes, four per
//        LDS.128, each half-warp its own row (the shipping kernel's scheme without padding, tails or staging);
//   G16  the same with 16-bit codes, eight per LDS.128;
//   GU   32-column rows (one row per warp instruction, warp-uniform 16-bit codes);
//   RED  A values in registers, B tile in shared memory, atomicAdd(double) per update (ATOMS.CAST.SPIN.64 loop on sm_100a);
//   RMW  A values in registers, B tile in shared memory, warp-private rows, plain LDS.64 + DADD + STS.64.
// Prints updates (one FP64 add into one element of B) per clock per SM at the measured SM clock (nvidia-smi max 1965 MHz
// assumed unless SM_MHZ is set).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o saso_update_rate saso_update_rate.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int THREADS = 512, ROWS_A = 512, PITCH = 17, NS = 32, REPS = 64;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ double lds64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ double sgn(double v, uint32_t code) {
    return __hiloint2double(__double2hiint(v) ^ (int)(code & 0x80000000u), __double2loint(v));
}

// MODE 0: G32, 1: G16, 2: GU
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) k_gather(double* out, int iters) {
    extern __shared__ __align__(128) uint8_t smem[];
    double* A = reinterpret_cast<double*>(smem);                                  // ROWS_A x PITCH(*2 for GU)
    constexpr int PW = MODE == 2 ? 33 : PITCH;
    uint32_t* codes = reinterpret_cast<uint32_t*>(smem + ROWS_A * PW * 8);       // per warp: NS slots x 2 halves x 16 bytes
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < ROWS_A * PW; i += THREADS) A[i] = 1.0 + i;
    uint32_t x = 12345u + tid * 747796405u + blockIdx.x;
    for (int i = tid; i < 16 * NS * 2 * 4 * 2; i += THREADS) {
        x = x * 1664525u + 1013904223u;
        uint32_t r0 = (x >> 8) % ROWS_A, r1 = (x >> 19) % ROWS_A;
        if (MODE == 0) codes[i] = (r0 * PW) | (x & 0x80000000u);
        else codes[i] = ((r0 * PW) | ((x & 1u) << 15)) | (((r1 * PW) | ((x & 2u) << 14)) << 16);
    }
    __syncthreads();
    double acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) acc[s] = 0.0;
    const int half = MODE == 2 ? 0 : lane >> 4, col = MODE == 2 ? lane : (lane & 15);
    uint32_t lanebase;
    asm volatile("mov.u32 %0, %1;" : "=r"(lanebase) : "r"(smem_u32(A) + col * 8));      // opaque: keeps the address one LEA
    const uint32_t recs0 = smem_u32(codes) + (warp * NS * 2 * 2 + half) * 16;      // 2 KB of records per warp
    for (int it = 0; it < iters; ++it) {
        const uint32_t recs = recs0 + (it & 7) * 128;      // the record addresses change every pass: ptxas cannot hoist the loads
        if (MODE == 0) {
#pragma unroll
            for (int s = 0; s < NS; s += 16) {      // 4 records = 16 codes of 16 consecutive slots (no padding in this probe)
                uint32_t c[16]; double v[16];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(c[4 * q]), "=r"(c[4 * q + 1]), "=r"(c[4 * q + 2]), "=r"(c[4 * q + 3])
                                 : "r"(recs + (s / 4 + q) * 32) : "memory");
#pragma unroll
                for (int q = 0; q < 16; ++q) v[q] = lds64(lanebase + (c[q] << 3));
#pragma unroll
                for (int q = 0; q < 16; ++q) acc[s + q] += sgn(v[q], c[q]);
            }
        } else {
#pragma unroll
            for (int s = 0; s < NS; s += 16) {      // 2 records = 16 sixteen-bit codes
                uint32_t c[8]; double v[16];
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(c[4 * q]), "=r"(c[4 * q + 1]), "=r"(c[4 * q + 2]), "=r"(c[4 * q + 3])
                                 : "r"(recs + (s / 8 + q) * 32) : "memory");
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    v[2 * q] = lds64(lanebase + ((c[q] & 0x7fffu) << 3));
                    v[2 * q + 1] = lds64(lanebase + ((c[q] >> 13) & 0x3fff8u));
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    acc[s + 2 * q] += sgn(v[2 * q], c[q] << 16);
                    acc[s + 2 * q + 1] += sgn(v[2 * q + 1], c[q]);
                }
            }
        }
    }
    double t = 0;
#pragma unroll
    for (int s = 0; s < NS; ++s) t += acc[s];
    out[blockIdx.x * THREADS + tid] = t;
}

// MODE 0: RED (atomicAdd), 1: RMW (warp-private rows)
template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) k_smemB(double* out, int iters) {
    extern __shared__ __align__(128) uint8_t smem[];
    double* B = reinterpret_cast<double*>(smem);                                  // 1024 rows x 16
    uint32_t* rows = reinterpret_cast<uint32_t*>(smem + 1024 * 16 * 8);          // per warp 64 row ids per half
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, half = lane >> 4, col = lane & 15;
    for (int i = tid; i < 1024 * 16; i += THREADS) B[i] = 0.0;
    uint32_t x = 999u + tid * 747796405u + blockIdx.x;
    for (int i = tid; i < 16 * 2 * 64; i += THREADS) {
        x = x * 1664525u + 1013904223u;
        const int w = i / 128;
        rows[i] = MODE == 0 ? (x >> 8) % 1024 : (uint32_t)(w * 64 + ((x >> 8) % 64));
    }
    __syncthreads();
    double a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 1.0 + tid + j;
    const uint32_t* myrows = rows + (warp * 2 + half) * 64;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t r = myrows[(j + it) & 63];
            double* p = &B[r * 16 + col];
            if (MODE == 0) atomicAdd(p, a[j & 7]);
            else { *reinterpret_cast<volatile double*>(p) = *reinterpret_cast<volatile double*>(p) + a[j & 7]; __syncwarp(); }
        }
    }
    __syncthreads();
    out[blockIdx.x * THREADS + tid] = B[tid];
}

template <typename K>
static void run(const char* name, K kern, size_t smem, int iters, double upd_per_thread_iter, double mhz) {
    double* out; cudaMalloc(&out, 148 * THREADS * 8);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    kern<<<148, THREADS, smem>>>(out, 2);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); kern<<<148, THREADS, smem>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const cudaError_t e = cudaDeviceSynchronize();
    const double upd = upd_per_thread_iter * iters * THREADS;          // per SM
    printf("%-4s %8.3f ms  %7.2f updates/clk/SM  (%s)\n", name, best, upd / (best * 1e-3 * mhz * 1e6), cudaGetErrorString(e));
    cudaFree(out);
}

int main() {
    const double mhz = getenv("SM_MHZ") ? atof(getenv("SM_MHZ")) : 1965.0;
    const int iters = 20000;
    run("G32", k_gather<0>, ROWS_A * PITCH * 8 + 16 * NS * 2 * 16 * 2, iters, NS, mhz);
    run("G16", k_gather<1>, ROWS_A * PITCH * 8 + 16 * NS * 2 * 16 * 2, iters, NS, mhz);
    run("GU", k_gather<2>, ROWS_A * 33 * 8 + 16 * NS * 2 * 16 * 2, iters, NS, mhz);
    run("RED", k_smemB<0>, 1024 * 16 * 8 + 16 * 2 * 64 * 4, iters / 8, 32, mhz);
    run("RMW", k_smemB<1>, 1024 * 16 * 8 + 16 * 2 * 64 * 4, iters / 8, 32, mhz);
    (void) REPS;
    return 0;
}
