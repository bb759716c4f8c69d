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

// Table loads of the two-step programs: the dispatch guarantees that their tables sit in shared memory, so
// say so (a generic load of a shared address takes the long-latency path through L1TEX address translation).
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float2 lds_f2(const float2* p)
{
    float2 v;
    asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"((unsigned)__cvta_generic_to_shared(p)));
    return v;
}
#define DFT_LDS(ptr) ::dft::lds_f2(ptr)
#else
#define DFT_LDS(ptr) (*(ptr))
#endif

template <int RA, int RB>
DFT_HD void reg_fft(float2 (&v)[RA * RB])
{
    if constexpr (RB == 1) dft_reg<RA>(v);
    else dft_pfa<RA, RB>(v);
}

// Strided-axis complex FFT of one shared-memory tile buf[n*P + j] (n < N rows, TM columns), results to
// global rows q*inner (+ column), natural order; `inverse` reverses the output index (IDFT = DFT read
// backwards; the four-step passes instead read their input reversed and come here with inverse = false).
// Columns j >= ncol of a partial tile are computed on whatever the tile holds and dropped.
template <int R1A, int R1B, int R2A, int R2B, int TM, bool FOUR, class Ctx>
DFT_HD void cols_two_step(float2* buf, int P, const float2* tw /* w_N^j */, float2* dst, long long inner, int ncol, float scale,
                          bool inverse, const float2* tw4 /* four-step: w_Ntot^j, else null */, int n1, Ctx& ctx,
                          const KParams& kp, bool split /* rows go to their owning ranks (KParams::split_n) */,
                          long long split_img, long long split_col0)
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
            v[a] = cmul(p[a * P], DFT_LDS(twp));            // (the dispatch guarantees this table sits in shared memory)
        });
        reg_fft<R2A, R2B>(v);
        if (FOUR && tw4) {
            // first pass of a four-step FFT (compiled in only for lengths that occur as four-step factors): output row q = k1 + R1*k2 of slab n1 is multiplied by w_Ntot^(n1*q)
            const float2* t4 = tw4 + n1 * k1;
            const int t4s = n1 * R1;
            static_for<0, R2>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                v[k2] = cmul(v[k2], DFT_LDT(t4 + k2 * t4s));
            });
        }
        // column of this work item in the output: j, or (second pass over a padded workspace, KParams::unflat_in) workspace
        // row k2 / column c unfolded again; padding columns are dropped
        long long jo = j;
        bool ok = j < ncol;
        if (kp.unflat_in && !split) {
            const long long jj = split_col0 + j;
            const long long kk = jj / kp.unflat_in, c = jj - kk * kp.unflat_in;
            ok = ok && c < kp.unflat_w;
            jo = kk * kp.unflat_w + c;
        }
        if (j < ncol && split) {
            static_for<0, R2>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                const int q = k1 + R1 * k2;
                const int qq = inverse ? (q == 0 ? 0 : N - q) : q;
                float2* d = split_elem_ptr(kp, split_img, qq, split_col0 + j);
                if (d) *d = float2{v[k2].x * scale, v[k2].y * scale};
            });
        } else if (ok) {
            // output row q = k1 + R1*k2, or (N - q) mod N for the inverse
            // one pointer walks down (or, for the inverse, up) the output rows
            const long long step = inverse ? -(long long)R1 * inner : (long long)R1 * inner;
            float2* d = dst + jo + (inverse ? (long long)(N - k1) * inner : (long long)k1 * inner);
            if (inverse && k1 == 0) {                    // (N - 0) mod N = 0: row 0 stays, the rest walks up from N - R1
                dst[jo] = float2{v[0].x * scale, v[0].y * scale};
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


// ---- rows ---------------------------------------------------------------------------------------
// M = R1*R2 packed complex points per row, G rows per tile at pitch P.  T[a*R1 + k1] = w_M^(a*k1) is the
// transposed inter-step twiddle table (consecutive work items of step B read consecutive entries).
//
// Step A: one work item per (row, n2), 32 per row (n2 < R2 <= 32 active) so that no half-warp straddles two
// rows; DFT over n1 in registers; all loads, barrier, all stores -- which lets the result land at pitch
// PA != P (the inverse re-pitches 721 -> 720 so that step B's 128-bit loads stay aligned).
// Step B: one work item per (row, k1): R2 contiguous values by 128-bit loads (R2/2 odd: conflict free),
// times T, DFT over n2; results Z[k1 + R1*k2] are handed to `emit(row, k1, v)`.
template <int R1A, int R1B, int R2A, int R2B, class Ctx, class Emit>
DFT_HD void rows_two_step(float2* buf, int nrows_tile, int P, int PA, const float2* T, bool barrier_before_emit, Ctx& ctx, Emit&& emit)
{
    constexpr int R1 = R1A * R1B, R2 = R2A * R2B;
    static_assert(R2 <= 32 && R2 % 2 == 0 && (R2 / 2) % 2 == 1, "step B reads R2/2 (odd) 16-byte units per line");
    // Work items past the end redo the last one (same values to the same places): no divergence and no
    // conditionally-live register arrays across the barriers.
    // step A (one round when the group has 32 threads per row)
    for (int i0 = 0; i0 < nrows_tile * 32; i0 += ctx.nthr) {
        const int i = i0 + ctx.tid;
        const int r = (i >> 5) < nrows_tile ? (i >> 5) : nrows_tile - 1;
        const int n2 = (i & 31) < R2 ? (i & 31) : R2 - 1;
        float2 v[R1];
        const float2* p = buf + r * P + n2;
        static_for<0, R1>([&](auto ac) { constexpr int a = decltype(ac)::value; v[a] = p[a * R2]; });
        reg_fft<R1A, R1B>(v);
        ctx.sync();
        float2* q = buf + r * PA + n2;
        static_for<0, R1>([&](auto cc) { constexpr int c = decltype(cc)::value; q[c * R2] = v[c]; });
        ctx.sync();
    }
    // step B
    for (int i0 = 0; i0 < nrows_tile * R1; i0 += ctx.nthr) {
        const int i = (i0 + ctx.tid) < nrows_tile * R1 ? (i0 + ctx.tid) : nrows_tile * R1 - 1;
        const int r = i / R1, k1 = i - r * R1;
        float2 v[R2];
        const float4* p4 = reinterpret_cast<const float4*>(buf + r * PA + k1 * R2);
        const float2* tw = T + k1;
        static_for<0, R2 / 2>([&](auto jc) {
            constexpr int j = decltype(jc)::value;
            const float4 q = p4[j];
            if constexpr (j == 0) v[0] = float2{q.x, q.y};
            else v[2 * j] = cmul(float2{q.x, q.y}, DFT_LDS(tw + (2 * j) * R1));
            v[2 * j + 1] = cmul(float2{q.z, q.w}, DFT_LDS(tw + (2 * j + 1) * R1));
        });
        reg_fft<R2A, R2B>(v);
        if (barrier_before_emit) ctx.sync();            // emit overwrites the tile: every line must be in registers
        emit(r, k1, v);
        if (barrier_before_emit) ctx.sync();
    }
}

// forward R2C rows: packed rows in buf (pitch M), results X[0..M] of each row to global `o` (row stride M+1).
// twU[k] = w_W^k.
template <int R1A, int R1B, int R2A, int R2B, class Ctx>
DFT_HD void rows_r2c_two_step(float2* buf, int nrows, int P, const float2* T, const float2* twU, float2* o, Ctx& ctx)
{
    constexpr int R1 = R1A * R1B, R2 = R2A * R2B, M = R1 * R2, Mo = M + 1;
    rows_two_step<R1A, R1B, R2A, R2B>(buf, nrows, P, P, T, true, ctx, [&](int r, int k1, const float2(&v)[R2]) {
        float2* q = buf + r * P + k1;                   // natural order: Z[k1 + R1*k2]
        static_for<0, R2>([&](auto kc) { constexpr int k2 = decltype(kc)::value; q[k2 * R1] = v[k2]; });
    });
    // real-FFT split: each thread keeps its frequencies k (and M-k) and walks down the rows
    for (int k = ctx.tid; k <= M / 2; k += ctx.nthr) {
        const int kk = M - k;
        const float2 w = DFT_LDS(twU + k);
        const float2* za = buf + k;
        const float2* zb = buf + (k == 0 ? 0 : kk);
        float2* oa = o + k;
        float2* ob = o + kk;
        for (int r = 0; r < nrows; ++r, za += P, zb += P, oa += Mo, ob += Mo) {
            const float2 z1 = *za, z2 = *zb;
            const float sx = z1.x + z2.x, sy = z1.y - z2.y;   // z1 + conj z2
            const float dx = z1.x - z2.x, dy = z1.y + z2.y;   // z1 - conj z2
            const float tx = 0.5f * (w.x * dx - w.y * dy), ty = 0.5f * (w.x * dy + w.y * dx);
            *oa = float2{0.5f * sx + ty, 0.5f * sy - tx};
            if (kk != k) *ob = float2{0.5f * sx - ty, -0.5f * sy - tx};
        }
    }
    ctx.sync();
}

// inverse C2R rows: one-sided rows X[0..M] in buf (pitch P = M+1), real rows of W = 2M floats to global `o`.
// twU[k] = w_W^k.  The inverse transform runs as conj(FFT(conj Z')).
template <int R1A, int R1B, int R2A, int R2B, class Ctx>
DFT_HD void rows_c2r_two_step(float2* buf, int nrows, int P, const float2* T, const float2* twU, float2* o, float scale, Ctx& ctx)
{
    constexpr int R1 = R1A * R1B, R2 = R2A * R2B, M = R1 * R2;
    // Hermitian merge, in place, stored conjugated and scaled: Z'[k] = (a + conj b) + i conj(w)(a - conj b)
    for (int k = ctx.tid; k <= M / 2; k += ctx.nthr) {
        const int kk = M - k;
        const float2 w = DFT_LDS(twU + k);
        float2* pa = buf + k;
        float2* pb = buf + kk;
        for (int r = 0; r < nrows; ++r, pa += P, pb += P) {
            const float2 a = *pa, b = *pb;
            if (k == 0) {
                *pa = float2{scale * (a.x + b.x), -scale * (a.x - b.x)};
                continue;
            }
            const float sx = a.x + b.x, sy = a.y - b.y;
            const float dx = a.x - b.x, dy = a.y + b.y;
            const float ux = w.x * dx + w.y * dy, uy = w.x * dy - w.y * dx;
            *pa = float2{scale * (sx - uy), -scale * (sy + ux)};
            if (kk != k) *pb = float2{scale * (sx + uy), -scale * (-sy + ux)};
        }
    }
    ctx.sync();
    rows_two_step<R1A, R1B, R2A, R2B>(buf, nrows, P, M, T, false, ctx, [&](int r, int k1, const float2(&v)[R2]) {
        float2* q = o + (long long)r * M + k1;          // x[2n], x[2n+1] = Re, -Im of the transform at n = k1 + R1*k2
        static_for<0, R2>([&](auto kc) { constexpr int k2 = decltype(kc)::value; q[k2 * R1] = float2{v[k2].x, -v[k2].y}; });
    });
    ctx.sync();
}

}  // namespace dft
