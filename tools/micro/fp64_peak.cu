// Microbenchmark: register-resident DMMA.8x8x4 and DFMA issue rates on one B200 (the FP64 roofline denominators).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu ; run: ./fp64_peak
#include <cstdio>
#include <cuda_runtime.h>
template <int NACC>
__global__ void __launch_bounds__(256) dmma_loop(double* out, int iters) {
    double c[NACC][2];
    for (int i = 0; i < NACC; ++i) { c[i][0] = 0; c[i][1] = 0; }
    double a = threadIdx.x * 1e-3, b = threadIdx.x * 2e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
__global__ void __launch_bounds__(256) dfma_loop(double* out, int iters) {
    double c[NACC];
    for (int i = 0; i < NACC; ++i) c[i] = i;
    double a = 1.0 + threadIdx.x * 1e-9, b = threadIdx.x * 2e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i) c[i] = fma(c[i], a, b);
    }
    double s = 0;
    for (int i = 0; i < NACC; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    double* out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(double));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int ctas = 1; ctas <= 4; ctas *= 2) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            dmma_loop<16><<<148 * ctas, 256>>>(out, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fl = 2.0 * 256 * 16 * (double) iters * 8 * 148 * ctas;
            if (rep) printf("DMMA.8x8x4  %d CTA/SM x 8 warps: %.2f ms  %.2f TFLOP/s\n", ctas, ms, fl / ms * 1e-9);
            cudaEventRecord(e0);
            dfma_loop<16><<<148 * ctas, 256>>>(out, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
            fl = 2.0 * 16 * (double) iters * 256 * 148 * ctas;
            if (rep) printf("DFMA        %d CTA/SM x 8 warps: %.2f ms  %.2f TFLOP/s\n", ctas, ms, fl / ms * 1e-9);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
