// Throughput of packed fp32 (FFMA2 / FADD2) against scalar FFMA / FADD on sm_100a: cycles per warp-instruction with
// enough independent chains and warps to saturate the pipe.  nvcc -arch=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { float2 v = make_float2(a, b); return *reinterpret_cast<u64*>(&v); }
template <int MODE>
__global__ void k(float* out, long long* cyc, float s)
{
    float a[16];
    u64 p[8];
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x + i;
    for (int i = 0; i < 8; ++i) p[i] = pk(a[2 * i], a[2 * i + 1]);
    const u64 ss = pk(s, s);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < 512; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, a[i]);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %1, %2, %1;" : "+l"(p[i]) : "l"(p[i]), "l"(ss));
        } else if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = a[i] + s;
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("add.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(ss));
        }
    }
    long long t1 = clock64();
    float r = 0;
    for (int i = 0; i < 16; ++i) r += a[i];
    for (int i = 0; i < 8; ++i) { float2 v = *reinterpret_cast<float2*>(&p[i]); r += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main()
{
    float* o; long long* c; cudaMalloc(&o, 1 << 24); cudaMalloc(&c, 8);
    const char* nm[4] = {"FFMA  x16 (16 fma)", "FFMA2 x8  (16 fma)", "FADD  x16", "FADD2 x8"};
    for (int warps = 4; warps <= 32; warps *= 2)
        for (int m = 0; m < 4; ++m) {
            long long h = 0;
            if (m == 0) k<0><<<148, warps * 32>>>(o, c, 1.0001f);
            if (m == 1) k<1><<<148, warps * 32>>>(o, c, 1.0001f);
            if (m == 2) k<2><<<148, warps * 32>>>(o, c, 1.0001f);
            if (m == 3) k<3><<<148, warps * 32>>>(o, c, 1.0001f);
            cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
            const double flops_per_sm_clk = (double)warps * 32 * 512 * 16 / h;      // fp32 lane-ops per clock per SM
            printf("warps/SM %2d  %-20s cycles %8lld  lane-ops per clk per SM %.1f\n", warps, nm[m], h, flops_per_sm_clk);
        }
    return 0;
}
