// dft_phases.cuh -- the per-tile phases of the shared-memory FFT engine.
//
// A tile (one or more images / rows / a column chunk) sits in shared memory as complex
// float2 values with pitch P.  Every phase is written as "thread tid of nthr does its share";
// phases are separated by a CTA barrier in the kernels (dft_kernels.cu).  The same functions
// compile for the host (tests/emu) where the "threads" of a phase run one after another --
// legal because no phase has an intra-phase dependency between threads.
//
// All FFT steps are forward (w = exp(-2*pi*i/n)), in place, decimation in frequency; the
// digit-reversed result order and the inverse transform's index reversal are resolved by the
// perm tables when the result leaves shared memory.
#pragma once
#include <math.h>

#include "dft_butterflies.cuh"
#include "dft_types.h"

#if defined(__CUDA_ARCH__)
#define DFT_LDG(ptr) __ldg(ptr)     // tensor data: always global memory
#else
#define DFT_LDG(ptr) (*(ptr))
#endif
#define DFT_LDT(ptr) (*(ptr))       // twiddle / permutation tables: global or shared memory (generic load)

namespace dft {

// ---- small integer helpers ---------------------------------------------------------------------
DFT_HD unsigned magic_of(unsigned d) { return d <= 1u ? 0u : (unsigned)(((1ull << 32) + d - 1) / d); }
DFT_HD int fdiv(int x, unsigned magic)   // floor(x / d) for the d that produced `magic`; exact for x*d < 2^32
{
    if (magic == 0u) return x;
#if defined(__CUDA_ARCH__)
    return (int)__umulhi((unsigned)x, magic);
#else
    return (int)(((unsigned long long)(unsigned)x * magic) >> 32);
#endif
}
// skewed shared-memory position of logical element p of a power-of-two row (pad != 0)
DFT_HD int phys(int p, int pad) { return pad ? p + (p >> 4) + (p >> 8) : p; }

// Column-split ("packed for all-to-all") address of element (global row R, column k); see KParams::pack_n.
DFT_HD float2* packed_ptr(const KParams& p, float2* out, int R, int k, int Mo)
{
    const int b = R / p.pack_hn, rl = R - b * p.pack_hn;
    // fully unrolled with constant indices: a run-time index into p.pack_c0 would push the kernel's
    // private KParams copy into local memory
    int c0 = p.pack_c0[0], cn = p.pack_c0[1] - p.pack_c0[0], pitch = p.pack_pitch[0];
    float2* peer = p.pack_peer[0];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int t = 1; t < 8; ++t) {
        if (t < p.pack_n && k >= p.pack_c0[t]) {
            c0 = p.pack_c0[t];
            cn = p.pack_c0[t + 1] - p.pack_c0[t];
            peer = p.pack_peer[t];
            pitch = p.pack_pitch[t];
        }
    }
    if (p.pack_peer_on)   // straight into the owning rank's slab [image][pack_H][pitch >= cn]
        return peer + ((long long)b * p.pack_H + p.pack_h0 + rl) * pitch + (k - c0);
    return out + (long long)b * p.pack_hn * Mo + (long long)p.pack_hn * c0 + (long long)rl * cn + (k - c0);
}

// Row-split address of element (pass row qp, pass column jj) of image o_ of a strided-axis pass (see KParams::split_n).
DFT_HD float2* split_row_ptr(const KParams& p, long long o_, int q);
DFT_HD float2* split_elem_ptr(const KParams& p, long long o_, int qp, long long jj)
{
    int q = qp;
    long long c = jj;
    if (p.split_n2 > 1) {
        const int k2 = (int)(jj / p.split_inner);
        c = jj - (long long)k2 * p.split_inner;
        q = qp * p.split_n2 + k2;
        if (p.unflat_w && c >= p.unflat_w) return nullptr;       // padding column of the workspace: nothing to store
    }
    return split_row_ptr(p, o_, q) + c;
}
// ... and of its row q (image rows), at this rank's first column
DFT_HD float2* split_row_ptr(const KParams& p, long long o_, int q)
{
    int r0 = p.split_r0[0], rn = p.split_r0[1] - p.split_r0[0];
    float2* peer = p.split_peer[0];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int t = 1; t < 8; ++t) {
        if (t < p.split_n && q >= p.split_r0[t]) {
            r0 = p.split_r0[t];
            rn = p.split_r0[t + 1] - p.split_r0[t];
            peer = p.split_peer[t];
        }
    }
    return peer + (o_ * rn + (q - r0)) * p.split_pitch + p.split_c0;
}

// ---- one radix-R step over all lines of the tile -------------------------------------------
// ROWS (COLS=false): line l, logical element e at buf[l*pitch + phys(e)].
// COLS (COLS=true) : line = (img, col) with ncols columns per image; element e at
//                    buf[img*img_stride + e*pitch + col]; consecutive threads take
//                    consecutive columns (bank-conflict free, stride 1).
// With pad != 0 the radix schedule is r_top,16,16,... so that within one butterfly the skew is
// linear in the element index (see dft_planner.h).
template <int R, bool COLS>
DFT_HD void fft_step(float2* buf, int nlines, int pitch, int N, int L, const float2* tw, int ncols,
                     int img_stride, unsigned mmag, unsigned bmag, unsigned lmag, int pad, int ord, int tid, int nthr,
                     const float2* tw_t = nullptr /* rows: this step's transposed twiddles, [(c-1)*m + b] */)
{
    const int m = L / R;         // element stride inside the butterfly
    const int bpl = N / R;       // butterflies per line
    const int tws = N / L;       // twiddle table stride: w_L^x = w_N^(x*tws)
    // rows: work items per line, padded to whole half-warps when the planner asked for it (LineFFT::ord)
    const int bplp = (!COLS && (ord & 2)) ? ((bpl + 15) & ~15) : bpl;
    const int total = nlines * bplp;
    const int es_rows = pad ? m + (m >> 4) + (m >> 8) : m;
    for (int t = tid; t < total; t += nthr) {
        int line, u;
        if (COLS) {
            u = fdiv(t, lmag);           // lmag: magic of nlines
            line = t - u * nlines;
        } else {
            line = fdiv(t, bmag);
            u = t - line * bplp;
            if (u >= bpl) continue;
        }
        int blk, b;
        if (m == 1) {
            blk = u;
            b = 0;
        } else if (!COLS && (ord & 1)) {     // consecutive items walk the blocks (mmag = magic of N/L)
            const int nblk = N / L;
            b = fdiv(u, mmag);
            blk = u - b * nblk;
        } else {
            blk = fdiv(u, mmag);
            b = u - blk * m;
        }
        const int e0 = blk * L + b;
        float2* p;
        int es;
        if (COLS) {
            int base = line;
            if (ncols != nlines) {  // more than one image in the tile
                const int img = line / ncols;
                base = img * img_stride + (line - img * ncols);
            }
            p = buf + base + e0 * pitch;
            es = m * pitch;
        } else {
            p = buf + line * pitch + phys(e0, pad);
            es = es_rows;
        }
        float2 v[R];
        static_for<0, R>([&](auto ac) {
            constexpr int a = decltype(ac)::value;
            v[a] = p[a * es];
        });
        dft_reg<R>(v);
        if (m > 1 && !COLS && tw_t) {
            const float2* twb = tw_t + b;
            static_for<1, R>([&](auto cc) {
                constexpr int c = decltype(cc)::value;
                v[c] = cmul(v[c], DFT_LDT(twb + (c - 1) * m));
            });
        } else if (m > 1) {
            const float2* twb = tw + (long long)b * tws;
            const int tstep = b * tws;
            static_for<1, R>([&](auto cc) {
                constexpr int c = decltype(cc)::value;
                const float2 w = DFT_LDT(twb + (c - 1) * tstep);
                v[c] = cmul(v[c], w);
            });
        }
        static_for<0, R>([&](auto cc) {
            constexpr int c = decltype(cc)::value;
            p[c * es] = v[c];
        });
    }
}

#if defined(__CUDACC__)
#define DFT_NOINLINE_HD __host__ __device__ __noinline__
#else
#define DFT_NOINLINE_HD
#endif
template <int R, bool COLS>
DFT_NOINLINE_HD void fft_step_big(float2* buf, int nlines, int pitch, int N, int L, const float2* tw, int ncols, int img_stride,
                                  unsigned mmag, unsigned bmag, unsigned lmag, int pad, int ord, int tid, int nthr, const float2* tw_t)
{
    fft_step<R, COLS>(buf, nlines, pitch, N, L, tw, ncols, img_stride, mmag, bmag, lmag, pad, ord, tid, nthr, tw_t);
}

// ---- one radix-R step for ANY prime R (used for prime factors above 31) ----------------------------------------------
// No register butterfly exists for these, so every output element is one work item: Y[c] = w_L^(b c) * sum_a x[a] w_R^(a c),
// R shared-memory reads and complex multiply-adds per output (O(N R) per line instead of the naive kernels' O(N^2) from
// global memory).  The step cannot run in place that way, so it writes into `scr`, a second copy of the tile's
// footprint; the caller copies it back after a barrier.  tw = w_N^j, j in [0, N).
template <bool COLS>
DFT_HD void fft_step_generic(int R, const float2* buf, float2* scr, int nlines, int pitch, int N, int L, const float2* tw, int ncols,
                             int img_stride, int tid, int nthr)
{
    const int m = L / R;
    const int tws = N / L, twr = N / R;
    const long long total = (long long)nlines * N;
    for (long long t = tid; t < total; t += nthr) {
        int line, e;
        if (COLS) {
            e = (int)(t / nlines);
            line = (int)(t - (long long)e * nlines);
        } else {
            line = (int)(t / N);
            e = (int)(t - (long long)line * N);
        }
        const int blk = e / L, r = e - blk * L;
        const int c = r / m, b = r - c * m;
        const int e0 = blk * L + b;
        int base, es;
        if (COLS) {
            base = line;
            if (ncols != nlines) {
                const int img = line / ncols;
                base = img * img_stride + (line - img * ncols);
            }
            base += e0 * pitch;
            es = m * pitch;
        } else {
            base = line * pitch + e0;
            es = m;
        }
        const float2* x = buf + base;
        // blocked summation (partial sums of 32 terms, then the sum of those): the rounding error grows like
        // sqrt(32) + sqrt(R / 32) ulps instead of sqrt(R), which keeps the round trip of a 4099-point prime inside 1e-6
        float ax = 0.f, ay = 0.f;
        int idx = 0;
        const int step = c * twr;                       // (a c mod R) * N/R, advanced incrementally
        for (int a0 = 0; a0 < R; a0 += 32) {
            float bx = 0.f, by = 0.f;
            const int a1 = a0 + 32 < R ? a0 + 32 : R;
            for (int a = a0; a < a1; ++a) {
                const float2 w = DFT_LDT(tw + idx), v = x[a * es];
                bx = fmaf(v.x, w.x, fmaf(-v.y, w.y, bx));
                by = fmaf(v.x, w.y, fmaf(v.y, w.x, by));
                idx += step;
                if (idx >= N) idx -= N;
            }
            ax += bx;
            ay += by;
        }
        if (m > 1) {
            const float2 w = DFT_LDT(tw + b * c * tws);   // b c < L, so b c tws < N
            const float tx = ax * w.x - ay * w.y;
            ay = ax * w.y + ay * w.x;
            ax = tx;
        }
        scr[base + c * es] = float2{ax, ay};
    }
}

DFT_HD bool radix_has_butterfly(int r)
{
    return r == 2 || r == 3 || r == 4 || r == 5 || r == 6 || r == 7 || r == 8 || r == 9 || r == 10 || r == 11 || r == 13 || r == 16 ||
           r == 17 || r == 19 || r == 23 || r == 29 || r == 31;
}

template <bool COLS>
DFT_HD void fft_step_dispatch(int radix, float2* buf, int nlines, int pitch, int N, int L,
                              const float2* tw, int ncols, int img_stride, unsigned mmag, unsigned bmag,
                              unsigned lmag, int pad, int ord, int tid, int nthr, const float2* tw_t = nullptr)
{
    switch (radix) {
#define DFT_CASE(R)                                                                     \
    case R:                                                                             \
        fft_step<R, COLS>(buf, nlines, pitch, N, L, tw, ncols, img_stride, mmag, bmag, lmag, pad, ord, tid, nthr, tw_t); \
        break;
// the large prime radices need more registers than the kernels' launch bounds leave: kept out of line, so that their
// spills stay inside them and the common radices keep the register allocation they had without them
#define DFT_CASE_BIG(R)                                                                 \
    case R:                                                                             \
        fft_step_big<R, COLS>(buf, nlines, pitch, N, L, tw, ncols, img_stride, mmag, bmag, lmag, pad, ord, tid, nthr, tw_t); \
        break;
        DFT_CASE(2)
        DFT_CASE(3)
        DFT_CASE(4)
        DFT_CASE(5)
        DFT_CASE(6)
        DFT_CASE(7)
        DFT_CASE(8)
        DFT_CASE(9)
        DFT_CASE(10)
        DFT_CASE(11)
        DFT_CASE(13)
        DFT_CASE(16)
        DFT_CASE_BIG(17)
        DFT_CASE_BIG(19)
        DFT_CASE_BIG(23)
        DFT_CASE_BIG(29)
        DFT_CASE_BIG(31)
#undef DFT_CASE
#undef DFT_CASE_BIG
        default:
            break;
    }
}

// ---- forward real-FFT split ("untangle") ----------------------------------------------------
// x real of length W = 2M, z[n] = x[2n] + i x[2n+1], Z = FFT_M(z):
//   X[k] = (Z[k] + conj Z[M-k])/2 - (i/2) w_W^k (Z[k] - conj Z[M-k]),   k = 0..M  (indices mod M)
// In 2-D the same formula holds after the column FFT with the partner taken at row (H-q) mod H.
DFT_HD float2 untangle(float2 z1, float2 z2, float2 w)
{
    // e = (z1 + conj z2)/2 ; o = (z1 - conj z2)/2 ; X = e - i*w*o
    const float ex = 0.5f * (z1.x + z2.x), ey = 0.5f * (z1.y - z2.y);
    const float ox = 0.5f * (z1.x - z2.x), oy = 0.5f * (z1.y + z2.y);
    // w*o
    const float tx = w.x * ox - w.y * oy, ty = w.x * oy + w.y * ox;
    // -i*(tx + i ty) = ty - i tx
    return float2{ex + ty, ey - tx};
}

// ---- inverse real-FFT merge (in place, per row) -----------------------------------------------
// From the one-sided row T[0..M] build Z'[k] = (T[k] + conj T[M-k]) + i conj(w_W^k)(T[k] - conj T[M-k]),
// k = 0..M-1, whose (unnormalised) inverse M-point FFT is W*x packed as x[2n] + i x[2n+1].
// The imaginary parts of T[0] and T[M] are ignored (C2R semantics).
DFT_HD void merge_rows(float2* buf, int nrows, int pitch, int M, const float2* tw_unt, int pad, int tid, int nthr)
{
    const int half = M / 2;                  // pairs k = 0..half  (k and M-k)
    const int per_row = half + 1;
    const int total = nrows * per_row;
    const unsigned pmag = magic_of((unsigned)per_row);
    for (int t = tid; t < total; t += nthr) {
        const int r = fdiv(t, pmag);
        const int k = t - r * per_row;
        float2* row = buf + r * pitch;
        if (k == 0) {
            const float a = row[0].x, b = row[phys(M, pad)].x;
            row[0] = float2{a + b, a - b};
            continue;
        }
        const int k2 = M - k;
        const int pk = phys(k, pad), pk2 = phys(k2, pad);
        const float2 a = row[pk], b = row[pk2];
        const float2 w = DFT_LDT(tw_unt + k);  // (c, s) = exp(-2 pi i k / W)
        // Z'[k]   = (a + conj b) + i*conj(w)*(a - conj b)
        // Z'[M-k] = (b + conj a) - i*w*(b - conj a)
        const float sx = a.x + b.x, sy = a.y - b.y;   // a + conj b
        const float dx = a.x - b.x, dy = a.y + b.y;   // a - conj b
        // conj(w)*(d) = (c - i s)(dx + i dy) = (c dx + s dy) + i(c dy - s dx)
        const float ux = w.x * dx + w.y * dy, uy = w.x * dy - w.y * dx;
        // i*u = (-uy, ux)
        row[pk] = float2{sx - uy, sy + ux};
        if (k2 != k) {
            // b + conj a = (sx, -sy) ; b - conj a = (-dx, dy)
            // w*(b - conj a) = (c + i s)(-dx + i dy) = (-c dx - s dy) + i(c dy - s dx) = (-ux, uy)
            // -i*(-ux + i uy) = (uy, ux)
            row[pk2] = float2{sx + uy, -sy + ux};
        }
    }
}

}  // namespace dft
