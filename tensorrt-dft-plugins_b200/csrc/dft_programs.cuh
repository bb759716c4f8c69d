// dft_programs.cuh -- per-mode tile programs: load -> FFT phases -> store.
//
// `Ctx` supplies tid / nthr / sync() and (device only) a bulk-async (TMA 1-D) loader.  The CUDA
// kernels instantiate the programs with a CTA context; tests/emu instantiates them with a
// one-thread host context (sync is a no-op), which checks all the index arithmetic on the CPU.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "dft_phases.cuh"

namespace dft {

// scratch_off != 0: the plan contains prime radices above 31; those steps go out of place into buf + scratch_off and are
// copied back (see fft_step_generic).  Compiled out of the kernels whose schedule is a compile-time constant.
template <class Ctx>
DFT_HD void run_line_fft(const LineFFT& f, bool cols, float2* buf, int nlines, int pitch,
                         const float2* tw, int ncols, int img_stride, Ctx& ctx, int pad = 0, const float2* tw_t = nullptr,
                         int scratch_off = 0)
{
    int L = f.n;
    int toff = 0;                                       // start of this step's block of tw_t
    const unsigned lmag = cols ? magic_of((unsigned)nlines) : 0u;   // nlines differs on the last, partial tile
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int s = 0; s < kMaxSteps; ++s) {
        if (s >= f.nsteps) break;
        const int r = f.radix[s];
        if (scratch_off != 0 && !radix_has_butterfly(r)) {
            float2* scr = buf + scratch_off;
            if (cols) fft_step_generic<true>(r, buf, scr, nlines, pitch, f.n, L, tw, ncols, img_stride, ctx.tid, ctx.nthr);
            else fft_step_generic<false>(r, buf, scr, nlines, pitch, f.n, L, tw, ncols, img_stride, ctx.tid, ctx.nthr);
            ctx.sync();
            for (int i = ctx.tid; i < scratch_off; i += ctx.nthr) buf[i] = scr[i];
            ctx.sync();
            toff += (r - 1) * (L / r);
            L /= r;
            continue;
        }
        if (cols)
            fft_step_dispatch<true>(r, buf, nlines, pitch, f.n, L, tw, ncols, img_stride, f.mmag[s], f.bmag[s], lmag, 0, 0,
                                    ctx.tid, ctx.nthr);
        else
            fft_step_dispatch<false>(r, buf, nlines, pitch, f.n, L, tw, ncols, img_stride, f.mmag[s], f.bmag[s], lmag, pad,
                                     f.ord[s], ctx.tid, ctx.nthr, tw_t ? tw_t + toff : nullptr);
        ctx.sync();
        toff += (r - 1) * (L / r);
        L /= r;
    }
}

// Contiguous global -> shared copy of `nfloats` floats; ends with the data visible to the CTA.
template <class Ctx>
DFT_HD void load_contig(Ctx& ctx, float* dst, const float* src, long long nfloats)
{
    if (ctx.bulk_load(dst, src, nfloats * 4)) return;  // device: cp.async.bulk + mbarrier wait
    const uintptr_t a = (uintptr_t)src;
    if ((a & 15) == 0 && (nfloats & 3) == 0) {
        const float4* s4 = (const float4*)src;
        float4* d4 = (float4*)dst;
        for (long long i = ctx.tid; i < nfloats / 4; i += ctx.nthr) d4[i] = DFT_LDG(s4 + i);
    } else if ((a & 7) == 0 && (nfloats & 1) == 0) {
        const float2* s2 = (const float2*)src;
        float2* d2 = (float2*)dst;
        for (long long i = ctx.tid; i < nfloats / 2; i += ctx.nthr) d2[i] = DFT_LDG(s2 + i);
    } else {
        for (long long i = ctx.tid; i < nfloats; i += ctx.nthr) dst[i] = DFT_LDG(src + i);
    }
    ctx.sync();
}

// ---- MODE_F2D: fused rfft2 of G images per tile ---------------------------------------------
template <class Ctx, bool LOAD = true>
DFT_HD void program_f2d(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int H = p.H, M = p.M, Mo = p.M + 1;
    const long long first = tile * p.G;
    const int nimg = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    if (LOAD) load_contig(ctx, (float*)buf, in + first * H * p.W, (long long)nimg * H * p.W);
    run_line_fft(p.row, false, buf, nimg * H, M, p.tw_row, 0, 0, ctx, 0, p.tw_row_t, p.scratch_off);
    run_line_fft(p.col, true, buf, nimg * M, M, p.tw_col, M, H * M, ctx, 0, nullptr, p.scratch_off);
    // 2-D real-FFT split, two outputs per item: X[q][k] and X[(H-q)%H][M-k] share the pair
    // Y[q][k], Y[(H-q)%H][(M-k)%M].  k = M/2 (M even) is produced twice with identical values.
    float2* o = (float2*)out + first * H * Mo;
    const int KH = M / 2 + 1;
    const int per_img = H * KH;
    const int total = nimg * per_img;
    const unsigned imag = magic_of((unsigned)per_img), kmag = magic_of((unsigned)KH);
    for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
        const int img = fdiv(idx, imag);
        const int rem = idx - img * per_img;
        const int q = fdiv(rem, kmag);
        const int k = rem - q * KH;
        const int kk = M - k;
        const int q2 = (q == 0) ? 0 : H - q;
        const float2* ib = buf + img * H * M;
        const float2 z1 = ib[DFT_LDT(p.perm_col + q) * M + DFT_LDT(p.perm_row + k)];
        const float2 z2 = ib[DFT_LDT(p.perm_col + q2) * M + DFT_LDT(p.perm_row + (k == 0 ? 0 : kk))];
        const float2 w = DFT_LDT(p.tw_unt + k);
        const float sx = z1.x + z2.x, sy = z1.y - z2.y;
        const float dx = z1.x - z2.x, dy = z1.y + z2.y;
        const float tx = 0.5f * (w.x * dx - w.y * dy), ty = 0.5f * (w.x * dy + w.y * dx);
        float2* oi = o + (long long)img * H * Mo;
        oi[q * Mo + k] = float2{0.5f * sx + ty, 0.5f * sy - tx};
        oi[q2 * Mo + kk] = float2{0.5f * sx - ty, -0.5f * sy - tx};
    }
    ctx.sync();
}

// ---- MODE_I2D: fused irfft2 --------------------------------------------------------------------
template <class Ctx, bool LOAD = true>
DFT_HD void program_i2d(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int H = p.H, M = p.M, P = p.M + 1;
    const long long first = tile * p.G;
    const int nimg = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    if (LOAD) load_contig(ctx, (float*)buf, in + first * H * P * 2, (long long)nimg * H * P * 2);
    run_line_fft(p.col, true, buf, nimg * P, P, p.tw_col, P, H * P, ctx, 0, nullptr, p.scratch_off);
    merge_rows(buf, nimg * H, P, M, p.tw_unt, 0, ctx.tid, ctx.nthr);
    ctx.sync();
    run_line_fft(p.row, false, buf, nimg * H, P, p.tw_row, 0, 0, ctx, 0, p.tw_row_t, p.scratch_off);
    float2* o = (float2*)out + first * H * M;
    const int per_img = H * M;
    const int total = nimg * per_img;
    const float sc = p.scale;
    const unsigned imag = magic_of((unsigned)per_img), mmag = magic_of((unsigned)M);
    for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
        const int img = fdiv(idx, imag);
        const int rem = idx - img * per_img;
        const int r = fdiv(rem, mmag);
        const int n = rem - r * M;
        const float2 z = buf[(img * H + DFT_LDT(p.perm_col + r)) * P + DFT_LDT(p.perm_row + n)];
        o[idx] = float2{z.x * sc, z.y * sc};
    }
    ctx.sync();
}

// ---- MODE_F1D: G rows per tile, W = 2M --------------------------------------------------------
// LOAD = false: the tile is already in `buf` (the pipelined kernel's producer warps put it there).
template <class Ctx, bool LOAD = true>
DFT_HD void program_f1d(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int M = p.M, Mo = p.M + 1, P = p.P;
    const long long first = tile * p.G;
    const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    if (!LOAD) {
    } else if (p.pad) {
        // power-of-two rows: skewed placement (phys), so no bulk copy; 8-byte loads when aligned
        const float* src = in + first * p.W;
        const unsigned mg = magic_of((unsigned)M);
        if ((((uintptr_t)src) & 7) == 0) {
            const float2* s2 = (const float2*)src;
            for (int idx = ctx.tid; idx < nrows * M; idx += ctx.nthr) {
                const int r = fdiv(idx, mg), n = idx - r * M;
                buf[r * P + phys(n, 1)] = DFT_LDG(s2 + idx);
            }
        } else {
            for (int idx = ctx.tid; idx < nrows * M; idx += ctx.nthr) {
                const int r = fdiv(idx, mg), n = idx - r * M;
                buf[r * P + phys(n, 1)] = float2{DFT_LDG(src + 2 * idx), DFT_LDG(src + 2 * idx + 1)};
            }
        }
        ctx.sync();
    } else {
        load_contig(ctx, (float*)buf, in + first * p.W, (long long)nrows * p.W);
    }
    run_line_fft(p.row, false, buf, nrows, P, p.tw_row, 0, 0, ctx, p.pad, p.tw_row_t, p.scratch_off);
    // real-FFT split, two outputs per item: X[k] and X[M-k] come from the same pair Z[k], Z[M-k]
    float2* o = (float2*)out + first * Mo;
    const int per_row = M / 2 + 1;
    // KParams::pack_bulk: results go to the group's staging rows (natural order) and leave as bulk copies below
    float2* stg = nullptr;
    if (p.pack_bulk) {
        stg = ctx.stage_out();
        ctx.wait_bulk_stores_read();                      // the previous tile's copies have read the staging rows
        ctx.sync();
    }
    const int total = nrows * per_row;
    const unsigned omg = magic_of((unsigned)per_row);
    for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
        const int r = fdiv(idx, omg);
        const int k = idx - r * per_row;
        const int kk = M - k;
        const float2* rb = buf + r * P;
        const float2 z1 = rb[DFT_LDT(p.perm_row + k)];
        const float2 z2 = rb[DFT_LDT(p.perm_row + (k == 0 ? 0 : kk))];
        const float2 w = DFT_LDT(p.tw_unt + k);
        const float sx = z1.x + z2.x, sy = z1.y - z2.y;   // z1 + conj z2
        const float dx = z1.x - z2.x, dy = z1.y + z2.y;   // z1 - conj z2
        const float tx = 0.5f * (w.x * dx - w.y * dy), ty = 0.5f * (w.x * dy + w.y * dx);
        const float2 x1 = float2{0.5f * sx + ty, 0.5f * sy - tx};
        const float2 x2 = float2{0.5f * sx - ty, -0.5f * sy - tx};
        if (stg) {
            float2* srow = stg + r * p.pack_mo_pad;
            srow[k] = x1;
            if (kk != k) srow[kk] = x2;
        } else if (p.pack_n == 0) {
            float2* orow = o + (long long)r * Mo;
            orow[k] = x1;
            if (kk != k) orow[kk] = x2;
        } else {
            float2* ob = (float2*)out;
            const int R = (int)first + r;
            *packed_ptr(p, ob, R, k, Mo) = x1;
            if (kk != k) *packed_ptr(p, ob, R, kk, Mo) = x2;
        }
    }
    if (stg) {
        // one bulk copy per (row, destination): columns [c0, c0 + cn) of the row into the owner's slab.  An odd cn is
        // rounded up: the extra element lands in the slab's padding column (pitch even, > cn) and comes from the staging
        // row's padding.
        ctx.fence_async_shared();
        ctx.sync();
        if (ctx.tid == 0) {
            for (int r = 0; r < nrows; ++r) {
                const int R = (int)first + r, b = R / p.pack_hn, rl = R - b * p.pack_hn;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
                for (int t = 0; t < 8; ++t) {
                    if (t >= p.pack_n) break;
                    const int c0 = p.pack_c0[t], cn = p.pack_c0[t + 1] - c0;
                    if (cn <= 0) continue;
                    float2* dst = p.pack_peer[t] + ((long long)b * p.pack_H + p.pack_h0 + rl) * p.pack_pitch[t];
                    ctx.bulk_store(dst, stg + r * p.pack_mo_pad + c0, (unsigned)((cn + 1) / 2 * 16));
                }
            }
            ctx.commit_bulk_stores();
        }
        return;                                           // (the next tile's first barrier follows its own loads)
    }
    ctx.sync();
}

// ---- MODE_I1D ------------------------------------------------------------------------------------
template <class Ctx, bool LOAD = true>
DFT_HD void program_i1d(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int M = p.M, Mo = p.M + 1, P = p.P;
    const long long first = tile * p.G;
    const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    if (!LOAD) {
    } else if (p.pad) {
        const float2* src = (const float2*)in + first * Mo;
        const unsigned mg = magic_of((unsigned)Mo);
        for (int idx = ctx.tid; idx < nrows * Mo; idx += ctx.nthr) {
            const int r = fdiv(idx, mg), k = idx - r * Mo;
            buf[r * P + phys(k, 1)] = DFT_LDG(src + idx);
        }
        ctx.sync();
    } else {
        load_contig(ctx, (float*)buf, in + first * Mo * 2, (long long)nrows * Mo * 2);
    }
    merge_rows(buf, nrows, P, M, p.tw_unt, p.pad, ctx.tid, ctx.nthr);
    ctx.sync();
    run_line_fft(p.row, false, buf, nrows, P, p.tw_row, 0, 0, ctx, p.pad, p.tw_row_t, p.scratch_off);
    float2* o = (float2*)out + first * M;
    const int total = nrows * M;
    const float sc = p.scale;
    const unsigned mg2 = magic_of((unsigned)M);
    for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
        const int r = fdiv(idx, mg2);
        const int n = idx - r * M;
        const float2 z = buf[r * P + DFT_LDT(p.perm_row + n)];
        o[idx] = float2{z.x * sc, z.y * sc};
    }
    ctx.sync();
}

// ---- MODE_F1D_ODD: odd W, rows as zero-imaginary complex lines of length W (= p.M) ---------------
template <class Ctx>
DFT_HD void program_f1d_odd(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int W = p.W, Mo = p.W / 2 + 1;
    const long long first = tile * p.G;
    const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    const float* src = in + first * W;
    for (int idx = ctx.tid; idx < nrows * W; idx += ctx.nthr) buf[idx] = float2{DFT_LDG(src + idx), 0.f};
    ctx.sync();
    run_line_fft(p.row, false, buf, nrows, W, p.tw_row, 0, 0, ctx, 0, nullptr, p.scratch_off);
    float2* o = (float2*)out + first * Mo;
    for (int idx = ctx.tid; idx < nrows * Mo; idx += ctx.nthr) {
        const int r = idx / Mo;
        const int k = idx - r * Mo;
        o[idx] = buf[r * W + DFT_LDT(p.perm_row + k)];
    }
    ctx.sync();
}

// ---- MODE_I1D_ODD: Hermitian extension to W complex bins, full FFT, real part --------------------
template <class Ctx>
DFT_HD void program_i1d_odd(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int W = p.W, Mo = p.W / 2 + 1;
    const long long first = tile * p.G;
    const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
    const float2* src = (const float2*)in + first * Mo;
    for (int idx = ctx.tid; idx < nrows * W; idx += ctx.nthr) {
        const int r = idx / W;
        const int n = idx - r * W;
        float2 v;
        if (n < Mo) {
            v = DFT_LDG(src + r * Mo + n);
        } else {
            v = DFT_LDG(src + r * Mo + (W - n));
            v.y = -v.y;
        }
        buf[idx] = v;
    }
    ctx.sync();
    run_line_fft(p.row, false, buf, nrows, W, p.tw_row, 0, 0, ctx, 0, nullptr, p.scratch_off);
    float* o = out + first * W;
    const float sc = p.scale;
    for (int idx = ctx.tid; idx < nrows * W; idx += ctx.nthr) {
        const int r = idx / W;
        const int n = idx - r * W;
        o[idx] = buf[r * W + DFT_LDT(p.perm_row + n)].x * sc;
    }
    ctx.sync();
}

// ---- MODE_C2C_COLS: data[outer][N][inner]; tile = (slab o, chunk c of G columns) ------------------
template <class Ctx, bool LOAD = true>
DFT_HD void program_cols(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    const int N = p.H, P = p.P;
    const long long o_ = tile / p.chunks;
    const long long c_ = tile - o_ * p.chunks;
    const long long col0 = c_ * p.G;
    const int ncol = (int)((p.inner - col0) < p.G ? (p.inner - col0) : p.G);
    // source rows: contiguous, or (four-step first pass) every n1split-th row starting at n1, optionally
    // index-reversed over the whole axis (inverse transform)
    const int n1s = p.n1split > 1 ? p.n1split : 1;
    const long long o_hi = o_ / n1s;
    const int n1 = (int)(o_ - o_hi * n1s);
    const long long ntot = (long long)n1s * N;
    const float2* img = (const float2*)in + o_hi * ntot * p.inner + col0;
    auto src_row = [&](int n) -> const float2* {
        long long row = n1 + (long long)n1s * n;
        if (p.rev_in) row = (row == 0) ? 0 : ntot - row;
        return img + row * p.inner;
    };
    const int total = N * ncol;
    // When the CTA size is a multiple of the tile width each thread keeps one column and walks
    // down the rows: no division in the loop, and the unrolled loads are all in flight together.
    const bool fixed_col = (ctx.nthr % ncol) == 0;
    const int rstep = ctx.nthr / ncol;
    const int j0 = ctx.tid % ncol, n0 = ctx.tid / ncol;
    if (!LOAD) {
    } else if (fixed_col) {
        float2* d0 = buf + n0 * P + j0;
        const int dstep = rstep * P;
#if defined(__CUDA_ARCH__)
#pragma unroll 8
#endif
        for (int n = n0; n < N; n += rstep, d0 += dstep) *d0 = DFT_LDG(src_row(n) + j0);
    } else {
        for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
            const int n = idx / ncol;
            const int j = idx - n * ncol;
            buf[n * P + j] = DFT_LDG(src_row(n) + j);
        }
    }
    if (LOAD) ctx.sync();
    run_line_fft(p.col, true, buf, ncol, P, p.tw_col, ncol, 0, ctx, 0, nullptr, p.scratch_off);
    const long long orow = p.out_rowlen ? p.out_rowlen : p.inner;          // pitch of the rows this pass writes
    float2* dst = (float2*)out + o_ * N * orow + col0;
    const float sc = p.scale;
    const bool four = p.n1split > 1;
    if (p.unflat_in && !p.split_n) {
        // second pass over a padded workspace: tile column j' = (workspace row k2, column c); see KParams::unflat_in
        const long long n2 = p.inner / p.unflat_in;
        float2* img = (float2*)out + o_ * N * n2 * p.unflat_w;
        for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
            const int q = idx / ncol;
            const int j = idx - q * ncol;
            const long long jj = col0 + j;
            const long long k2 = jj / p.unflat_in, c = jj - k2 * p.unflat_in;
            if (c >= p.unflat_w) continue;
            const float2 z = buf[DFT_LDT(p.perm_col + q) * P + j];
            img[((long long)q * n2 + k2) * p.unflat_w + c] = float2{z.x * sc, z.y * sc};
        }
    } else if (p.split_n) {
        // rows go to the rank that owns them (never combined with the four-step first pass, which writes the workspace)
        for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
            const int q = idx / ncol;
            const int j = idx - q * ncol;
            const float2 z = buf[DFT_LDT(p.perm_col + q) * P + j];
            float2* d = split_elem_ptr(p, o_, q, col0 + j);
            if (d) *d = float2{z.x * sc, z.y * sc};
        }
    } else if (fixed_col) {
        float2* d0 = dst + (long long)n0 * orow + j0;
        const long long dstep2 = (long long)rstep * orow;
        if (!four) {
#if defined(__CUDA_ARCH__)
#pragma unroll 4
#endif
            for (int q = n0; q < N; q += rstep, d0 += dstep2) {
                const float2 z = buf[DFT_LDT(p.perm_col + q) * P + j0];
                *d0 = float2{z.x * sc, z.y * sc};
            }
        } else {
            for (int q = n0; q < N; q += rstep, d0 += dstep2) {
                const float2 z = cmul(buf[DFT_LDT(p.perm_col + q) * P + j0], DFT_LDT(p.tw_four + (long long)n1 * q));
                *d0 = float2{z.x * sc, z.y * sc};
            }
        }
    } else {
        for (int idx = ctx.tid; idx < total; idx += ctx.nthr) {
            const int q = idx / ncol;
            const int j = idx - q * ncol;
            float2 z = buf[DFT_LDT(p.perm_col + q) * P + j];
            if (four) z = cmul(z, DFT_LDT(p.tw_four + (long long)n1 * q));
            dst[(long long)q * orow + j] = float2{z.x * sc, z.y * sc};
        }
    }
    ctx.sync();
}

template <class Ctx>
DFT_HD void run_tile(const KParams& p, long long tile, float2* buf, const float* in, float* out, Ctx& ctx)
{
    switch (p.mode) {
        case MODE_F2D: program_f2d(p, tile, buf, in, out, ctx); break;
        case MODE_I2D: program_i2d(p, tile, buf, in, out, ctx); break;
        case MODE_F1D: program_f1d(p, tile, buf, in, out, ctx); break;
        case MODE_I1D: program_i1d(p, tile, buf, in, out, ctx); break;
        case MODE_F1D_ODD: program_f1d_odd(p, tile, buf, in, out, ctx); break;
        case MODE_I1D_ODD: program_i1d_odd(p, tile, buf, in, out, ctx); break;
        case MODE_C2C_COLS: program_cols(p, tile, buf, in, out, ctx); break;
        default: break;
    }
}

// ---- naive O(N^2) passes: any length, used only when a length has a prime factor the radix set
// does not cover or a line does not fit in shared memory.  One output element per index.
// tw = w_n^j table of the axis length; sign < 0 forward, > 0 inverse.
DFT_HD void naive_rows_r2c_elem(const KParams& p, long long idx, const float* in, float* out)
{
    const int W = p.W, Mo = p.W / 2 + 1;
    const long long r = idx / Mo;
    const int k = (int)(idx - r * Mo);
    const float* x = in + r * W;
    double sr = 0.0, si = 0.0;
    for (int n = 0; n < W; ++n) {
        const float2 w = DFT_LDT(p.tw_row + (int)(((long long)n * k) % W));
        const double v = (double)DFT_LDG(x + n);
        sr += v * w.x;
        si += v * w.y;
    }
    ((float2*)out)[idx] = float2{(float)sr, (float)si};
}

DFT_HD void naive_rows_c2r_elem(const KParams& p, long long idx, const float* in, float* out)
{
    const int W = p.W, Mo = p.W / 2 + 1;
    const long long r = idx / W;
    const int n = (int)(idx - r * W);
    const float2* X = (const float2*)in + r * Mo;
    double s = (double)DFT_LDG(X).x;
    for (int k = 1; k < Mo; ++k) {
        const float2 w = DFT_LDT(p.tw_row + (int)(((long long)n * k) % W));  // (c, -s); inverse uses conj
        const float2 v = DFT_LDG(X + k);
        if (2 * k == W)
            s += (double)v.x * w.x;
        else
            s += 2.0 * ((double)v.x * w.x + (double)v.y * w.y);
    }
    out[idx] = (float)(s * (double)p.scale);
}

// data[outer][N][inner], out-of-place; p.W != 0 selects the inverse sign.
DFT_HD void naive_cols_elem(const KParams& p, long long idx, const float* in, float* out)
{
    const int N = p.H;
    const long long slab = (long long)N * p.inner;
    const long long o_ = idx / slab;
    const long long rem = idx - o_ * slab;
    const int q = (int)(rem / p.inner);
    const long long i = rem - (long long)q * p.inner;
    const float2* src = (const float2*)in + o_ * slab + i;
    const double sg = p.W ? -1.0 : 1.0;
    double sr = 0.0, si = 0.0;
    for (int n = 0; n < N; ++n) {
        const float2 w = DFT_LDT(p.tw_col + (int)(((long long)n * q) % N));
        const float2 v = DFT_LDG(src + (long long)n * p.inner);
        const double wy = sg * w.y;
        sr += (double)v.x * w.x - (double)v.y * wy;
        si += (double)v.x * wy + (double)v.y * w.x;
    }
    ((float2*)out)[idx] = float2{(float)(sr * p.scale), (float)(si * p.scale)};
}

}  // namespace dft
