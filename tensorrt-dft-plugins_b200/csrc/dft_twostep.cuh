// dft_twostep.cuh -- two-step, register-resident line FFTs for registered lengths N = R1 * R2.
//
// The generic engine (dft_programs.cuh) makes one trip through shared memory per radix step (three for
// 720 = 10*9*8) with run-time index arithmetic.  Here a line is transformed in two steps, each holding a
// whole sub-transform in registers, every index a compile-time constant:
//
//   n = R2*n1 + n2,  k = k1 + R1*k2
//   A : one work item per (n2, column): DFT over n1 (length R1 = R1A*R1B, Good-Thomas inside), written
//       back in place
//   B : one work item per (k1, column): loads times w_N^(n2*k1), DFT over n2 (length R2 = R2A*R2B), result
//       k = k1 + R1*k2 goes straight from registers to global memory
//
// i.e. two shared-memory trips instead of three, about 2.5x fewer instructions per element.
#pragma once
#include "dft_butterflies.cuh"
#include "dft_phases.cuh"

namespace dft {

template <int RA, int RB>
DFT_HD void reg_fft(float2 (&v)[RA * RB])
{
    if constexpr (RB == 1) dft_reg<RA>(v);
    else dft_pfa<RA, RB>(v);
}

// Strided-axis complex FFT of one shared-memory tile buf[n*P + j] (n < N rows, TM columns), results to
// global rows q*inner (+ column), natural order; `inverse` reverses the output index (IDFT = DFT read
// backwards).  Columns j >= ncol of a partial tile are computed on whatever the tile holds and dropped.
template <int R1A, int R1B, int R2A, int R2B, int TM, class Ctx>
DFT_HD void cols_two_step(float2* buf, int P, const float2* tw /* w_N^j */, float2* dst, long long inner, int ncol, float scale,
                          bool inverse, Ctx& ctx)
{
    constexpr int R1 = R1A * R1B, R2 = R2A * R2B, N = R1 * R2;
    // step A
    for (int i = ctx.tid; i < R2 * TM; i += ctx.nthr) {
        const int n2 = i / TM, j = i - n2 * TM;
        float2* p = buf + n2 * P + j;
        float2 v[R1];
        static_for<0, R1>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * R2 * P]; });
        reg_fft<R1A, R1B>(v);
        static_for<0, R1>([&](auto cc) { constexpr int c = decltype(cc)::value; p[c * R2 * P] = v[c]; });
    }
    ctx.sync();
    // step B
    for (int i = ctx.tid; i < R1 * TM; i += ctx.nthr) {
        const int k1 = i / TM, j = i - k1 * TM;
        const float2* p = buf + k1 * R2 * P + j;
        float2 v[R2];
        // the inter-step twiddle w_N^(n2*k1) rides on the loads (each one is used at once: no extra live registers)
        v[0] = p[0];
        const float2* twp = tw;
        static_for<1, R2>([&](auto ac) {
            constexpr int a = decltype(ac)::value;
            twp += k1;
            v[a] = cmul(p[a * P], DFT_LDT(twp));
        });
        reg_fft<R2A, R2B>(v);
        if (j < ncol) {
            // output row q = k1 + R1*k2, or (N - q) mod N for the inverse
            // one pointer walks down (or, for the inverse, up) the output rows
            const long long step = inverse ? -(long long)R1 * inner : (long long)R1 * inner;
            float2* d = dst + j + (inverse ? (long long)(N - k1) * inner : (long long)k1 * inner);
            if (inverse && k1 == 0) {                    // (N - 0) mod N = 0: row 0 stays, the rest walks up from N - R1
                dst[j] = float2{v[0].x * scale, v[0].y * scale};
            } else {
                *d = float2{v[0].x * scale, v[0].y * scale};
            }
            static_for<1, R2>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                d += step;
                *d = float2{v[k2].x * scale, v[k2].y * scale};
            });
        }
    }
    ctx.sync();
}

}  // namespace dft
