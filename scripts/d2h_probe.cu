// D2H of the step's result block: DMA copies (1-D, chunked 2-D like ctr_step_host) against an SM-driven export kernel that
// stores straight into the mapped pinned host buffer.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o d2h_probe d2h_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void k_export(const double4* __restrict__ src, double4* __restrict__ dst, size_t pitch4, size_t w4, int rows) {
    // rows x w4 double4 elements, row pitch pitch4 (both sides)
    const size_t total = (size_t)rows * w4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / w4, c = i - r * w4;
        dst[r * pitch4 + c] = src[r * pitch4 + c];
    }
}

int main() {
    const size_t N = 131072, ROWS = 21, CH = 8, per = N / CH;
    const size_t bytes = ROWS * N * sizeof(double);
    double *dev, *host;
    CK(cudaMalloc(&dev, bytes));
    CK(cudaMemset(dev, 1, bytes));
    CK(cudaHostAlloc(&host, bytes, cudaHostAllocDefault));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    auto timeit = [&](const char* name, auto fn) {
        for (int i = 0; i < 3; ++i) fn();
        CK(cudaStreamSynchronize(st));
        float best = 1e9f, sum = 0;
        for (int i = 0; i < 10; ++i) {
            CK(cudaEventRecord(a, st)); fn(); CK(cudaEventRecord(b, st)); CK(cudaStreamSynchronize(st));
            float ms; CK(cudaEventElapsedTime(&ms, a, b)); sum += ms; if (ms < best) best = ms;
        }
        printf("%-58s mean %.3f ms  best %.3f ms  %.1f GB/s\n", name, sum / 10, best, bytes / (sum / 10 * 1e-3) / 1e9);
    };
    timeit("cudaMemcpyAsync 1-D, 22 MB", [&] { CK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st)); });
    timeit("8 x cudaMemcpy2DAsync (21 rows x 128 KB)", [&] {
        for (size_t c = 0; c < CH; ++c)
            CK(cudaMemcpy2DAsync(host + c * per, N * 8, dev + c * per, N * 8, per * 8, ROWS, cudaMemcpyDeviceToHost, st));
    });
    uint8_t *dev8, *host8;
    CK(cudaMalloc(&dev8, 2 * N));
    CK(cudaHostAlloc(&host8, 2 * N, cudaHostAllocDefault));
    timeit("8 x (2-D f64 block + 2-D u8 flags), as ctr_step_host does", [&] {
        for (size_t c = 0; c < CH; ++c) {
            CK(cudaMemcpy2DAsync(host + c * per, N * 8, dev + c * per, N * 8, per * 8, ROWS, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpy2DAsync(host8 + c * per, N, dev8 + c * per, N, per, 2, cudaMemcpyDeviceToHost, st));
        }
    });
    timeit("8 x 2-D f64 block with one more row (flags as a 22nd row)", [&] {
        for (size_t c = 0; c < CH; ++c)
            CK(cudaMemcpy2DAsync(host + c * per, N * 8, dev + c * per, N * 8, per * 8, ROWS, cudaMemcpyDeviceToHost, st));
    });
    for (int blocks : {8}) {
        for (int threads : {256}) {
            char name[96];
            snprintf(name, sizeof name, "8 x k_export<<<%d,%d>>> (SM stores to mapped host)", blocks, threads);
            timeit(name, [&] {
                for (size_t c = 0; c < CH; ++c)
                    k_export<<<blocks, threads, 0, st>>>((const double4*)(dev + c * per), (double4*)(host + c * per), N / 4, per / 4, (int)ROWS);
            });
        }
    }
    timeit("1 x k_export<<<148,1024>>> whole block", [&] {
        k_export<<<148, 1024, 0, st>>>((const double4*)dev, (double4*)host, N / 4, N / 4, (int)ROWS);
    });
    CK(cudaGetLastError());
    return 0;
}
