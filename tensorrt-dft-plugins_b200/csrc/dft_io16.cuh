// dft_io16.cuh -- 16-bit elements at a kernel's boundary (typed-I/O plans, trt_dft_plan_create_typed): conversions between
// a complex value held as two fp16 / bf16 numbers in one 32-bit register and float2.  Arithmetic stays fp32.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

namespace dft {

#if defined(__CUDACC__)

template <class IO>
struct Io16;
// fold(a, b, sg) = a + sg * b for complex values a, b held as two 16-bit numbers in one register, sg = +-1 in the same
// 16-bit format: widen a, then one mixed-precision FMA per component (fma.f32.f16 / .bf16 -> FHFMA on sm_100a: the 16-bit
// operand is read as it is, exactly as a conversion followed by an fp32 FMA would compute).  Four instructions per complex
// value where convert-everything-then-FFMA takes six (fp16) or eight (bf16); the fp32 kernel takes two.
template <>
struct Io16<__half> {
    typedef __half2 v2;
    static constexpr unsigned short kOne = 0x3C00, kMinusOne = 0xBC00;
    __device__ static __forceinline__ v2 down(float2 v) { return __floats2half2_rn(v.x, v.y); }
    __device__ static __forceinline__ float2 up(v2 v) { return __half22float2(v); }
    __device__ static __forceinline__ float2 fold(unsigned a, unsigned b, unsigned short sg)
    {
        const unsigned short alo = (unsigned short)(a & 0xffffu), ahi = (unsigned short)(a >> 16);
        const unsigned short blo = (unsigned short)(b & 0xffffu), bhi = (unsigned short)(b >> 16);
        float ax, ay, rx, ry;
        asm("cvt.f32.f16 %0, %1;" : "=f"(ax) : "h"(alo));
        asm("cvt.f32.f16 %0, %1;" : "=f"(ay) : "h"(ahi));
        asm("fma.rn.f32.f16 %0, %1, %2, %3;" : "=f"(rx) : "h"(blo), "h"(sg), "f"(ax));
        asm("fma.rn.f32.f16 %0, %1, %2, %3;" : "=f"(ry) : "h"(bhi), "h"(sg), "f"(ay));
        return float2{rx, ry};
    }
};
template <>
struct Io16<__nv_bfloat16> {
    typedef __nv_bfloat162 v2;
    static constexpr unsigned short kOne = 0x3F80, kMinusOne = 0xBF80;
    __device__ static __forceinline__ v2 down(float2 v) { return __floats2bfloat162_rn(v.x, v.y); }
    __device__ static __forceinline__ float2 up(v2 v)
    {
        const unsigned u = reinterpret_cast<const unsigned&>(v);
        return float2{__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u)};
    }
    __device__ static __forceinline__ float2 fold(unsigned a, unsigned b, unsigned short sg)
    {
        const unsigned short blo = (unsigned short)(b & 0xffffu), bhi = (unsigned short)(b >> 16);
        const float ax = __uint_as_float(a << 16), ay = __uint_as_float(a & 0xffff0000u);      // bf16 is the top half of an fp32
        float rx, ry;
        asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(rx) : "h"(blo), "h"(sg), "f"(ax));
        asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(ry) : "h"(bhi), "h"(sg), "f"(ay));
        return float2{rx, ry};
    }
};

#endif  // __CUDACC__

}  // namespace dft
