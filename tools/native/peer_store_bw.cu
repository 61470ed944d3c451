// Remote (peer, NVLink) store bandwidth from one GPU into another: 8-byte against 16-byte stores per lane, one stream
// of addresses against two (keys and flags of the record exchange).  nvcc -O3 -arch=sm_100a peer_store_bw.cu -o peer_store_bw
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void st8x2(unsigned long long *a, unsigned long long *b, size_t n) {     // two 8-byte streams (what xi_scatter does)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        a[i] = i;
        b[i] = ~i;
    }
}
__global__ void st16(ulonglong2 *a, size_t n) {                                     // one 16-byte stream
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        a[i] = make_ulonglong2(i, ~i);
}
__global__ void copy8x2(const unsigned long long *sa, const unsigned long long *sb, unsigned long long *a, unsigned long long *b, size_t n) {
    // local loads -> remote stores, four records per thread in flight (what a scatter does, minus the index arithmetic)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4 * stride) {
        unsigned long long k[4], f[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + u * stride < n) { k[u] = __ldg(sa + i + u * stride); f[u] = __ldg(sb + i + u * stride); }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + u * stride < n) { a[i + u * stride] = k[u]; b[i + u * stride] = f[u]; }
    }
}
int main() {
    int nd = 0;
    CK(cudaGetDeviceCount(&nd));
    if (nd < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t n = 4u << 20;   // 4 M records of 16 bytes = 64 MB
    CK(cudaSetDevice(1));
    void *remote; CK(cudaMalloc(&remote, n * 16));
    CK(cudaSetDevice(0));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    void *local; CK(cudaMalloc(&local, n * 16));
    void *src; CK(cudaMalloc(&src, n * 16)); CK(cudaMemset(src, 1, n * 16));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int where = 0; where < 2; ++where) {
        char *dst = (char *)(where ? remote : local);
        for (int ctas : {148, 592, 1184, 4736}) {
            float ms8 = 0, ms16 = 0;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaEventRecord(e0));
                st8x2<<<ctas, 256>>>((unsigned long long *)dst, (unsigned long long *)(dst + n * 8), n);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms8, e0, e1));
                CK(cudaEventRecord(e0));
                st16<<<ctas, 256>>>((ulonglong2 *)dst, n);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms16, e0, e1));
            }
            float msc = 0;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaEventRecord(e0));
                copy8x2<<<ctas, 256>>>((unsigned long long *)src, (unsigned long long *)((char *)src + n * 8), (unsigned long long *)dst,
                                       (unsigned long long *)(dst + n * 8), n);
                CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&msc, e0, e1));
            }
            printf("%s %5d CTAs: copy (local loads -> stores) %.1f us = %.0f GB/s written\n", where ? "remote" : "local ", ctas, msc * 1e3,
                   n * 16 / msc * 1e-6);
            printf("%s %5d CTAs: 2 x 8-byte streams %.1f us = %.0f GB/s, 16-byte stores %.1f us = %.0f GB/s\n", where ? "remote" : "local ",
                   ctas, ms8 * 1e3, n * 16 / ms8 * 1e-6, ms16 * 1e3, n * 16 / ms16 * 1e-6);
        }
    }
    return 0;
}
