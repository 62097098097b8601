// Host->device copy rates for a column-major slab: cudaMemcpy2DAsync vs one cudaMemcpyAsync per column piece.
#include <cstdio>
#include <chrono>
#include <cuda_runtime.h>
int main() {
    const size_t N = 2048, M = 1 << 20, kl = 32768;           // A is M x N doubles column-major; slab = kl rows
    double* h; cudaHostAlloc(&h, N * M * 8, cudaHostAllocDefault);
    double* d; cudaMalloc(&d, N * kl * 8 * 4);
    cudaStream_t s; cudaStreamCreate(&s);
    auto now = [] { return std::chrono::steady_clock::now(); };
    for (int rep = 0; rep < 2; ++rep) {
        auto t0 = now();
        cudaMemcpy2DAsync(d, kl * 8, h, M * 8, kl * 8, N, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s);
        double dt = std::chrono::duration<double>(now() - t0).count();
        printf("2D  width %zu KB x %zu: %.2f ms  %.1f GB/s\n", kl * 8 / 1024, N, dt * 1e3, N * kl * 8 / dt / 1e9);
        t0 = now();
        for (size_t j = 0; j < N; ++j) cudaMemcpyAsync(d + j * kl, h + j * M, kl * 8, cudaMemcpyHostToDevice, s);
        auto t1 = now(); cudaStreamSynchronize(s);
        dt = std::chrono::duration<double>(now() - t0).count();
        printf("1D x %zu pieces of %zu KB: %.2f ms  %.1f GB/s (issue %.2f ms)\n", N, kl * 8 / 1024, dt * 1e3, N * kl * 8 / dt / 1e9,
               std::chrono::duration<double>(t1 - t0).count() * 1e3);
        t0 = now();
        for (size_t j = 0; j < N; ++j) cudaMemcpyAsync(d + j * kl * 4, h + j * M, kl * 8 * 4, cudaMemcpyHostToDevice, s);
        t1 = now(); cudaStreamSynchronize(s);
        dt = std::chrono::duration<double>(now() - t0).count();
        printf("1D x %zu pieces of %zu KB: %.2f ms  %.1f GB/s (issue %.2f ms)\n", N, kl * 8 * 4 / 1024, dt * 1e3, N * kl * 8 * 4 / dt / 1e9,
               std::chrono::duration<double>(t1 - t0).count() * 1e3);
        t0 = now();
        cudaMemcpy2DAsync(d, kl * 8 * 4, h, M * 8, kl * 8 * 4, N, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s);
        dt = std::chrono::duration<double>(now() - t0).count();
        printf("2D  width %zu KB x %zu: %.2f ms  %.1f GB/s\n", kl * 8 * 4 / 1024, N, dt * 1e3, N * kl * 8 * 4 / dt / 1e9);
        t0 = now();
        cudaMemcpyAsync(d, h, N * kl * 8, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s);
        dt = std::chrono::duration<double>(now() - t0).count();
        printf("1D contiguous: %.2f ms  %.1f GB/s\n", dt * 1e3, N * kl * 8 / dt / 1e9);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
