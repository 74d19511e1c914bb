// H2D probe: how fast do the x|y halves of n Jacobian points (96-byte records, 64 useful bytes)
// reach HBM from pinned host memory?  1-D copy of everything, 2-D (strided) copy of the useful
// bytes, and a kernel reading the mapped host buffer directly.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/h2d_probe tools/h2d_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e = (x);                                                           \
        if (e != cudaSuccess) {                                                        \
            printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); \
            return 1;                                                                  \
        }                                                                              \
    } while (0)

// one thread per 16 bytes of output: 4 threads per point
__global__ void k_pull(const uint4 *__restrict__ host, uint4 *__restrict__ dst, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n * 4; i += stride) {
        size_t p = i >> 2, q = i & 3;
        dst[i] = host[p * 6 + q];
    }
}

int main(int argc, char **argv) {
    size_t n = argc > 1 ? (size_t)1 << atoi(argv[1]) : (size_t)1 << 24;
    void *h, *d;
    CK(cudaSetDeviceFlags(cudaDeviceMapHost));
    CK(cudaHostAlloc(&h, n * 96, cudaHostAllocMapped));
    for (size_t i = 0; i < n * 96 / 8; i++) ((uint64_t *)h)[i] = i * 0x9E3779B97F4A7C15ull;
    CK(cudaMalloc(&d, n * 96));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    float ms;
    printf("{\"n\": %zu", n);
    for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(a, st));
        CK(cudaMemcpyAsync(d, h, n * 96, cudaMemcpyHostToDevice, st));
        CK(cudaEventRecord(b, st));
        CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b));
    }
    printf(", \"copy1d_96_ms\": %.3f, \"copy1d_GBps\": %.2f", ms, n * 96 / ms / 1e6);
    for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(a, st));
        CK(cudaMemcpy2DAsync(d, 64, h, 96, 64, n, cudaMemcpyHostToDevice, st));
        CK(cudaEventRecord(b, st));
        CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b));
    }
    printf(", \"copy2d_64of96_ms\": %.3f, \"copy2d_useful_GBps\": %.2f", ms, n * 64 / ms / 1e6);
    void *hd;
    CK(cudaHostGetDevicePointer(&hd, h, 0));
    for (int blocks = 148; blocks <= 148 * 16; blocks *= 4) {
        for (int rep = 0; rep < 2; rep++) {
            CK(cudaEventRecord(a, st));
            k_pull<<<blocks, 256, 0, st>>>((const uint4 *)hd, (uint4 *)d, n);
            CK(cudaEventRecord(b, st));
            CK(cudaEventSynchronize(b));
            CK(cudaEventElapsedTime(&ms, a, b));
        }
        printf(", \"pull_%d_blocks_ms\": %.3f, \"pull_%d_useful_GBps\": %.2f", blocks, ms, blocks,
               n * 64 / ms / 1e6);
    }
    printf("}\n");
    return 0;
}
