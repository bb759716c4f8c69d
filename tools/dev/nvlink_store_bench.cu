// Store bandwidth into a peer GPU's memory over NVLink from a plain kernel: 8-byte vs 16-byte stores per thread, and
// different numbers of CTAs.  One process, two GPUs (cudaDeviceEnablePeerAccess).
//   nvcc -arch=sm_100a -O3 -o nvlink_store_bench nvlink_store_bench.cu && ./nvlink_store_bench
#include <cstdio>
#include <cuda_runtime.h>
template <typename T>
__global__ void copy_k(const T* __restrict__ src, T* __restrict__ dst, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
int main()
{
    int nd = 0; cudaGetDeviceCount(&nd);
    if (nd < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t bytes = 256u << 20;
    void *src, *peer, *local;
    cudaSetDevice(1); cudaMalloc(&peer, bytes);
    cudaSetDevice(0); cudaDeviceEnablePeerAccess(1, 0); cudaMalloc(&src, bytes); cudaMalloc(&local, bytes);
    cudaMemset(src, 1, bytes);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int target = 0; target < 2; ++target)
        for (int grid : {148, 296, 592, 1184})
            for (int w = 8; w <= 16; w *= 2) {
                void* dst = target ? peer : local;
                auto run = [&]() {
                    if (w == 8) copy_k<float2><<<grid, 512>>>((const float2*)src, (float2*)dst, bytes / 8);
                    else copy_k<float4><<<grid, 512>>>((const float4*)src, (float4*)dst, bytes / 16);
                };
                run(); cudaDeviceSynchronize();
                cudaEventRecord(a);
                for (int i = 0; i < 10; ++i) run();
                cudaEventRecord(b); cudaEventSynchronize(b);
                float ms; cudaEventElapsedTime(&ms, a, b);
                printf("%s  grid %4d x 512  %2d-byte stores  %7.1f GB/s\n", target ? "peer " : "local", grid, w, bytes * 10.0 / ms / 1e6);
            }
    return 0;
}
