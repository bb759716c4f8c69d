// dft_typed_io.cu -- fp16 / bf16 <-> fp32 conversion passes behind the typed-I/O plans (trt_dft_plan_create_typed) of
// shapes that have no kernel reading and writing 16-bit data itself: the fp32 transform runs through the workspace and
// these two passes sit either side of it.  Arithmetic is fp32 everywhere; only what crosses HBM at the plan's boundary
// is 16-bit.  Round to nearest even on the way out.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "dft_launch.h"

namespace {

template <class T>
__device__ __forceinline__ float widen1(T v);
template <>
__device__ __forceinline__ float widen1<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float widen1<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <class T>
__device__ __forceinline__ T narrow1(float v);
template <>
__device__ __forceinline__ __half narrow1<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 narrow1<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// eight elements per thread and trip (16 bytes of 16-bit data, 32 bytes of fp32) when both pointers allow it
template <class T>
__global__ void __launch_bounds__(256) widen_kernel(const T* __restrict__ src, float* __restrict__ dst, long long n, int vec)
{
    const long long tid = blockIdx.x * 256LL + threadIdx.x, nthr = (long long)gridDim.x * 256;
    long long done = 0;
    if (vec) {
        const long long n8 = n / 8;
        for (long long i = tid; i < n8; i += nthr) {
            const uint4 raw = __ldg(reinterpret_cast<const uint4*>(src) + i);
            const T* h = reinterpret_cast<const T*>(&raw);
            float4 a = float4{widen1<T>(h[0]), widen1<T>(h[1]), widen1<T>(h[2]), widen1<T>(h[3])};
            float4 b = float4{widen1<T>(h[4]), widen1<T>(h[5]), widen1<T>(h[6]), widen1<T>(h[7])};
            reinterpret_cast<float4*>(dst)[2 * i] = a;
            reinterpret_cast<float4*>(dst)[2 * i + 1] = b;
        }
        done = n8 * 8;
    }
    for (long long i = done + tid; i < n; i += nthr) dst[i] = widen1<T>(src[i]);
}

template <class T>
__global__ void __launch_bounds__(256) narrow_kernel(const float* __restrict__ src, T* __restrict__ dst, long long n, int vec)
{
    const long long tid = blockIdx.x * 256LL + threadIdx.x, nthr = (long long)gridDim.x * 256;
    long long done = 0;
    if (vec) {
        const long long n8 = n / 8;
        for (long long i = tid; i < n8; i += nthr) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(src) + 2 * i);
            const float4 b = __ldg(reinterpret_cast<const float4*>(src) + 2 * i + 1);
            uint4 raw;
            T* h = reinterpret_cast<T*>(&raw);
            h[0] = narrow1<T>(a.x), h[1] = narrow1<T>(a.y), h[2] = narrow1<T>(a.z), h[3] = narrow1<T>(a.w);
            h[4] = narrow1<T>(b.x), h[5] = narrow1<T>(b.y), h[6] = narrow1<T>(b.z), h[7] = narrow1<T>(b.w);
            reinterpret_cast<uint4*>(dst)[i] = raw;
        }
        done = n8 * 8;
    }
    for (long long i = done + tid; i < n; i += nthr) dst[i] = narrow1<T>(src[i]);
}

int grid_for(long long n)
{
    long long b = (n / 8 + 255) / 256;
    if (b > 148LL * 16) b = 148LL * 16;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace

namespace dft {

void launch_widen(const void* src, float* dst, long long n, int io_type, void* stream)
{
    if (n <= 0) return;
    const int vec = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
    if (io_type == 1) widen_kernel<__half><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((const __half*)src, dst, n, vec);
    else widen_kernel<__nv_bfloat16><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)src, dst, n, vec);
}

void launch_narrow(const float* src, void* dst, long long n, int io_type, void* stream)
{
    if (n <= 0) return;
    const int vec = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
    if (io_type == 1) narrow_kernel<__half><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(src, (__half*)dst, n, vec);
    else narrow_kernel<__nv_bfloat16><<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n, vec);
}

}  // namespace dft
