// dft_butterflies.cuh -- in-register DFT butterflies (forward sign, w = exp(-2*pi*i/R)).
//
// Every index is a compile-time constant (static_for + integral constants), so a float2 v[R]
// lives entirely in registers and every internal twiddle is an immediate.  Radices: the
// primes 2,3,5,7,11,13,17,19,23,29,31 (symmetric half-sum form); 6,10,12,15 by a twiddle-free Good-Thomas
// prime-factor split; 8,9,16 by one Cooley-Tukey split with constant twiddles.  The inverse transform never needs its own
// butterflies: IDFT(x)[n] = DFT(x)[(N-n) mod N], and the index reversal is folded into the
// permutation tables used when results are stored (see dft_planner.cpp).
#pragma once
#include "dft_types.h"

namespace dft {

template <int I>
struct IC {
    static constexpr int value = I;
};

template <int B, int E, class F>
DFT_HD void static_for(F&& f)
{
    if constexpr (B < E) {
        f(IC<B>{});
        static_for<B + 1, E>(f);
    }
}

// ---- compile-time cos/sin of 2*pi*j/n with exact octant reduction ---------------------------
#if defined(__CUDACC__)
#define DFT_CX __host__ __device__ constexpr
#else
#define DFT_CX constexpr
#endif
constexpr double kPi = 3.14159265358979323846264338327950288;

DFT_CX double cx_sin_poly(double x)  // |x| <= pi/4
{
    const double x2 = x * x;
    double term = x, sum = x;
    for (int k = 1; k <= 12; ++k) {
        term *= -x2 / double((2 * k) * (2 * k + 1));
        sum += term;
    }
    return sum;
}

DFT_CX double cx_cos_poly(double x)  // |x| <= pi/4
{
    const double x2 = x * x;
    double term = 1.0, sum = 1.0;
    for (int k = 1; k <= 12; ++k) {
        term *= -x2 / double((2 * k - 1) * (2 * k));
        sum += term;
    }
    return sum;
}

DFT_CX double cx_cos2pi(long long j, long long n)  // cos(2*pi*j/n)
{
    j %= n;
    if (j < 0) j += n;
    long long num = j, den = n;
    if (2 * num > den) num = den - num;  // cos(2pi f) = cos(2pi (1-f))
    bool neg = false;
    if (4 * num > den) {                 // f in (1/4,1/2]: cos = -cos(2pi (1/2-f))
        num = den - 2 * num;
        den = 2 * den;
        neg = true;
    }
    double r = 0.0;
    if (8 * num > den) {                 // f in (1/8,1/4]: cos = sin(2pi (1/4-f))
        r = cx_sin_poly(2.0 * kPi * double(den - 4 * num) / double(4 * den));
    } else {
        r = cx_cos_poly(2.0 * kPi * double(num) / double(den));
    }
    return neg ? -r : r;
}

DFT_CX double cx_sin2pi(long long j, long long n)  // sin(2*pi*j/n) = cos(2*pi*(j/n - 1/4))
{
    return cx_cos2pi(4 * j - n, 4 * n);
}

// ---- complex helpers ------------------------------------------------------------------------
// DFT_F32X2 (sm_100a): complex add / subtract as ONE packed instruction (add.f32x2 -> FADD2).  The packed form has the
// same lane throughput as two scalar adds (measured, tools/dev/ffma2_bench.cu) but takes one issue slot instead of two.
#if defined(__CUDA_ARCH__) && defined(DFT_F32X2)
__device__ __forceinline__ unsigned long long f2_bits(float2 a) { return (unsigned long long)__float_as_uint(a.x) | ((unsigned long long)__float_as_uint(a.y) << 32); }
__device__ __forceinline__ float2 bits_f2(unsigned long long r) { return float2{__uint_as_float((unsigned)r), __uint_as_float((unsigned)(r >> 32))}; }
__device__ __forceinline__ float2 cadd(float2 a, float2 b)
{
    unsigned long long r;
    asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 csub(float2 a, float2 b)
{
    unsigned long long r;
    asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(b)));
    return bits_f2(r);
}
// p + c * a with a real scalar c (both halves): one FFMA2 with a broadcast operand
__device__ __forceinline__ float2 cfma_s(float c, float2 a, float2 p)
{
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(float2{c, c})), "l"(f2_bits(p)));
    return bits_f2(r);
}
__device__ __forceinline__ float2 cmul_s(float c, float2 a)
{
    unsigned long long r;
    asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_bits(a)), "l"(f2_bits(float2{c, c})));
    return bits_f2(r);
}
#else
DFT_HD float2 cadd(float2 a, float2 b) { return float2{a.x + b.x, a.y + b.y}; }
DFT_HD float2 csub(float2 a, float2 b) { return float2{a.x - b.x, a.y - b.y}; }
DFT_HD float2 cfma_s(float c, float2 a, float2 p) { return float2{fmaf(c, a.x, p.x), fmaf(c, a.y, p.y)}; }
DFT_HD float2 cmul_s(float c, float2 a) { return float2{c * a.x, c * a.y}; }
#endif
DFT_HD float2 cmul(float2 a, float2 w) { return float2{a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x}; }
DFT_HD float2 cconj(float2 a) { return float2{a.x, -a.y}; }

// a * w_N^J  (w_N = exp(-2*pi*i/N)), with the trivial rotations resolved at compile time
template <int N, int J>
DFT_HD float2 mul_w(float2 a)
{
    constexpr int j = ((J % N) + N) % N;
    if constexpr (j == 0) {
        return a;
    } else if constexpr (4 * j == N) {
        return float2{a.y, -a.x};
    } else if constexpr (2 * j == N) {
        return float2{-a.x, -a.y};
    } else if constexpr (4 * j == 3 * N) {
        return float2{-a.y, a.x};
    } else if constexpr (8 * j == N) {
        constexpr float h = 0.70710678118654752440f;
        return float2{(a.x + a.y) * h, (a.y - a.x) * h};
    } else if constexpr (8 * j == 3 * N) {
        constexpr float h = 0.70710678118654752440f;
        return float2{(a.y - a.x) * h, -(a.x + a.y) * h};
    } else if constexpr (8 * j == 5 * N) {
        constexpr float h = 0.70710678118654752440f;
        return float2{-(a.x + a.y) * h, (a.x - a.y) * h};
    } else if constexpr (8 * j == 7 * N) {
        constexpr float h = 0.70710678118654752440f;
        return float2{(a.x - a.y) * h, (a.x + a.y) * h};
    } else {
        constexpr float c = float(cx_cos2pi(j, N));
        constexpr float s = float(-cx_sin2pi(j, N));
        return float2{a.x * c - a.y * s, a.x * s + a.y * c};
    }
}

// ---- prime radices: X_k = p_k - i q_k, X_{P-k} = p_k + i q_k ---------------------------------
template <int P>
DFT_HD void dft_prime(float2 (&v)[P])
{
    static_assert(P % 2 == 1 && P >= 3, "odd prime expected");
    constexpr int Hh = (P - 1) / 2;
    // a_j = v_j + v_{P-j};  br_j = -i (v_j - v_{P-j}) = (d.y, -d.x): kept rotated, so that X_k = p_k + qr_k and
    // X_{P-k} = p_k - qr_k are plain complex additions (packed on sm_100a, see cadd)
    float2 a[Hh], br[Hh];
    static_for<0, Hh>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        a[j] = cadd(v[j + 1], v[P - 1 - j]);
        br[j] = float2{v[j + 1].y - v[P - 1 - j].y, v[P - 1 - j].x - v[j + 1].x};
    });
    const float2 v0 = v[0];
    float2 s0 = v0;
    static_for<0, Hh>([&](auto jc) { s0 = cadd(s0, a[decltype(jc)::value]); });
    v[0] = s0;
    static_for<1, Hh + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        float2 p = v0;
        float2 q = float2{0.f, 0.f};
        static_for<1, Hh + 1>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            constexpr float c = float(cx_cos2pi((long long)j * k, P));
            constexpr float s = float(cx_sin2pi((long long)j * k, P));
            p = cfma_s(c, a[j - 1], p);
            if constexpr (j == 1) q = cmul_s(s, br[0]);
            else q = cfma_s(s, br[j - 1], q);
        });
        v[k] = cadd(p, q);
        v[P - k] = csub(p, q);
    });
}

template <int R>
DFT_HD void dft_reg(float2 (&v)[R]);

// One Cooley-Tukey split N = R1*R2: n = n1*R2 + n2, k = k1 + R1*k2.
template <int R1, int R2>
DFT_HD void dft_ct(float2 (&v)[R1 * R2])
{
    constexpr int N = R1 * R2;
    float2 t[N];  // t[k1*R2 + n2]
    static_for<0, R2>([&](auto n2c) {
        constexpr int n2 = decltype(n2c)::value;
        float2 s[R1];
        static_for<0, R1>([&](auto n1c) {
            constexpr int n1 = decltype(n1c)::value;
            s[n1] = v[n1 * R2 + n2];
        });
        dft_reg<R1>(s);
        static_for<0, R1>([&](auto k1c) {
            constexpr int k1 = decltype(k1c)::value;
            t[k1 * R2 + n2] = mul_w<N, n2 * k1>(s[k1]);
        });
    });
    static_for<0, R1>([&](auto k1c) {
        constexpr int k1 = decltype(k1c)::value;
        float2 s[R2];
        static_for<0, R2>([&](auto n2c) {
            constexpr int n2 = decltype(n2c)::value;
            s[n2] = t[k1 * R2 + n2];
        });
        dft_reg<R2>(s);
        static_for<0, R2>([&](auto k2c) {
            constexpr int k2 = decltype(k2c)::value;
            v[k1 + R1 * k2] = s[k2];
        });
    });
}

// Good-Thomas prime-factor split N = R1*R2 with gcd(R1,R2) = 1: a true R1 x R2 two-dimensional
// DFT, no twiddles.  n = (R2*n1 + R1*n2) mod N,  k = (R2*(R2^-1 mod R1)*k1 + R1*(R1^-1 mod R2)*k2) mod N.
DFT_CX int cx_modinv(int a, int m)
{
    a %= m;
    for (int x = 1; x < m; ++x)
        if ((a * x) % m == 1) return x;
    return 1;
}

template <int R1, int R2>
DFT_HD void dft_pfa(float2 (&v)[R1 * R2])
{
    constexpr int N = R1 * R2;
    constexpr int I2 = cx_modinv(R2, R1), I1 = cx_modinv(R1, R2);
    float2 t[N];  // t[k1*R2 + n2]
    static_for<0, R2>([&](auto n2c) {
        constexpr int n2 = decltype(n2c)::value;
        float2 s[R1];
        static_for<0, R1>([&](auto n1c) {
            constexpr int n1 = decltype(n1c)::value;
            s[n1] = v[(R2 * n1 + R1 * n2) % N];
        });
        dft_reg<R1>(s);
        static_for<0, R1>([&](auto k1c) {
            constexpr int k1 = decltype(k1c)::value;
            t[k1 * R2 + n2] = s[k1];
        });
    });
    static_for<0, R1>([&](auto k1c) {
        constexpr int k1 = decltype(k1c)::value;
        float2 s[R2];
        static_for<0, R2>([&](auto n2c) {
            constexpr int n2 = decltype(n2c)::value;
            s[n2] = t[k1 * R2 + n2];
        });
        dft_reg<R2>(s);
        static_for<0, R2>([&](auto k2c) {
            constexpr int k2 = decltype(k2c)::value;
            v[(R2 * I2 * k1 + R1 * I1 * k2) % N] = s[k2];
        });
    });
}

template <int R>
DFT_HD void dft_reg(float2 (&v)[R])
{
    if constexpr (R == 1) {
    } else if constexpr (R == 2) {
        const float2 a = v[0], b = v[1];
        v[0] = cadd(a, b);
        v[1] = csub(a, b);
    } else if constexpr (R == 4) {
        const float2 t0 = cadd(v[0], v[2]), t1 = csub(v[0], v[2]);
        const float2 t2 = cadd(v[1], v[3]), d = csub(v[1], v[3]);
        const float2 t3 = float2{d.y, -d.x};  // -i*(v1 - v3)
        v[0] = cadd(t0, t2);
        v[2] = csub(t0, t2);
        v[1] = cadd(t1, t3);
        v[3] = csub(t1, t3);
    } else if constexpr (R == 3 || R == 5 || R == 7 || R == 11 || R == 13 || R == 17 || R == 19 || R == 23 || R == 29 || R == 31) {
        dft_prime<R>(v);
    } else if constexpr (R == 6) {
        dft_pfa<2, 3>(v);
    } else if constexpr (R == 8) {
        dft_ct<2, 4>(v);
    } else if constexpr (R == 9) {
        dft_ct<3, 3>(v);
    } else if constexpr (R == 10) {
        dft_pfa<2, 5>(v);
    } else if constexpr (R == 12) {
        dft_pfa<3, 4>(v);
    } else if constexpr (R == 15) {
        dft_pfa<3, 5>(v);
    } else if constexpr (R == 16) {
        dft_ct<4, 4>(v);
    } else if constexpr (R == 32) {
        dft_ct<4, 8>(v);
    } else {
        static_assert(R == 1, "unsupported radix");
    }
}

}  // namespace dft
