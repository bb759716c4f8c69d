// dft_reg2d.cuh -- persistent fused rfft2 / irfft2 kernels that keep whole lines in registers.
//
// For images whose two FFT lengths are 2 * odd (M = W/2 = 2*Mh, H = 2*Hh; 90 = 2*45 for the 90 x 180
// AFNO grid) a length-M line is split by Good-Thomas into a radix-2 step and a length-Mh FFT:
//
//     n = (Mh*n1 + 2*n2) mod M,   k = (Mh*k1 + (Mh+1)*k2) mod M,
//     Z[k1,k2] = sum_n2 w_Mh^(n2 k2) * ( z[2*n2] + (-1)^k1 * z[(Mh + 2*n2) mod M] )
//
// Two threads share a line, one per parity k1 = p.  Each reads the whole line from shared memory
// (128-bit loads), forms its Mh sums or differences and runs the length-Mh FFT entirely in registers
// (a second Good-Thomas split RA x RB, no twiddles, every index a compile-time constant).  There is no
// exchange step between butterfly passes at all: an image makes ONE trip through shared memory between
// the row transform and the column transform (the transpose), where the pass-per-radix kernels of
// dft_pipe2d.cuh make four.  The real-FFT split pairs Z[k] with Z[M-k], which have the same parity, so it
// also stays inside the thread.
//
//   forward : TMA bulk load of the image -> rows (thread = (row, parity)): load, FFT, split, store
//             transposed [k][row] -> columns (thread = (column k, parity)): load, FFT, coalesced 8-byte
//             global stores.  The stage is handed back as soon as the column loads have landed.
//   inverse : TMA bulk load of the spectrum -> columns (thread = (k, parity)): load (index-reversed, which
//             turns the forward butterflies into the inverse transform), FFT, store transposed [k][row]
//             -> rows: parities sit in the two halves of a warp; Hermitian merge of the thread's own
//             parity, FFT, radix-2 step LAST (decimation in time) through one warp shuffle per value,
//             store the dense real image -> one TMA bulk store.
//
// Bank behaviour: a thread's line is contiguous (pitch M*8 = 720 bytes, 45 16-byte units: odd), lanes
// run across lines, so the 128-bit loads of a quarter warp hit 8 different 16-byte bank groups; the
// transposed stores have lanes along the contiguous index.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "dft_butterflies.cuh"
#include "dft_io16.cuh"
#include "dft_phases.cuh"
#include "dft_pipe2d.cuh"

namespace dft {

template <int H_, int W_, int RA_, int RB_, int CA_, int CB_, int NG_, int NS_, bool IOW_ = false, int PP_ = 0>
struct RegCfg {
    // PP: the two groups take turns -- a group starts the first half of an image (load + first transform) only
    // when the other group has moved on to its second half, so finished images leave at even intervals.
    static constexpr int PP = PP_;
    static_assert(PP_ == 0 || NG_ == 2, "ping-pong is between two groups");
    // IOW: one extra warp per compute group drives that group's bulk stores and the refills that follow
    // them (the compute warps then never wait on a store), at the price of a 128-register budget.
    static constexpr bool IOW = IOW_;
    static constexpr int H = H_, W = W_, M = W_ / 2, Mh = W_ / 4, Hh = H_ / 2, MO = W_ / 2 + 1;
    static constexpr int RA = RA_, RB = RB_, CA = CA_, CB = CB_, NG = NG_, NS = NS_;
    static_assert(W_ % 4 == 0 && Mh % 2 == 1 && H_ % 2 == 0 && Hh % 2 == 1, "both FFT lengths must be 2 * odd");
    static_assert(RA_ * RB_ == Mh && CA_ * CB_ == Hh, "bad factorisation");
    static constexpr int LINES = (H > MO ? H : MO);
    static constexpr int WPP = (LINES + 31) / 32;          // warps per parity
    static constexpr int GT = 2 * WPP * 32;                // threads per compute group
    static constexpr int THREADS = NG * GT + (IOW ? NG * 32 : 0);
    static constexpr int LCM = (NG * NS) / (NG % NS == 0 ? NS : (NS % NG == 0 ? NG : 1));   // NG, NS coprime or nested
    static constexpr int TILE_F = H * M * 8;               // real image == packed complex rows
    static constexpr int TILE_I = H * MO * 8;              // one-sided spectrum
    static_assert(TILE_F % 16 == 0 && TILE_I % 16 == 0, "bulk copies need 16-byte multiples");
    static_assert((M / 2) % 2 == 1 && (H / 2) % 2 == 1, "line pitch must be odd in 16-byte units");
    static constexpr int STAGE_F = (MO * H * 8 + 127) / 128 * 128;     // transposed [MO][H]
    static constexpr int PI = (H % 2 == 0) ? H + 1 : H;                // inverse transpose pitch: odd
    static constexpr int STAGE_I = (MO * PI * 8 + 127) / 128 * 128;
    // truncated inverse: the kept block [2*kh][kw] is loaded so that it ENDS at element TAIL of the stage (even: 16-byte
    // aligned start), which lets it be expanded in place into the spectrum tile [q][k] front to back
    static constexpr int TAIL = (MO * PI) / 2 * 2;
    static_assert(TAIL >= H * MO + 2 && TAIL * 8 <= STAGE_I, "kept block must fit behind the spectrum tile's expansion front");
    static constexpr int NPAIR = (Mh - 1) / 2;
    static constexpr int NTAB_I = 2 * (NPAIR + 1);         // inverse merge twiddles, per parity
    static constexpr int NTAB_F = 2 * (NPAIR + 1);         // forward split twiddles, per parity
    static constexpr int SMEM_F = NS * STAGE_F + NTAB_F * 8 + (NS + NG * NS) * 8 + 16;
    static constexpr int SMEM_I = NS * STAGE_I + NTAB_I * 8 + (NS + NG * NS) * 8 + 16;
};

#if defined(__CUDACC__)

namespace reg2d {

#ifdef DFT_REG2D_TRACE
__device__ long long* g_trace;   // [image slot][8] clock64 stamps of CTA 0, debugging only
#define DFT_TR(slot, i) do { if (blockIdx.x == 0 && (gt & 31) == 0 && warp % C::WPP == 0 && p == 0) reg2d::g_trace[(slot) * 8 + (i)] = clock64(); } while (0)
#else
#define DFT_TR(slot, i) do { } while (0)
#endif

// y[n2] = x[2*n2] + sg * x[(L2 + 2*n2) mod 2*L2] from a contiguous line of 2*L2 complex values (L2 odd)
template <int L2>
__device__ __forceinline__ void load_fold(float2 (&y)[L2], const float4* __restrict__ line, float sg)
{
    // 16-byte unit j holds x[2j], x[2j+1]; the odd partner of n2 sits in unit (n2 + (L2-1)/2) mod L2.
    // Walking n2 in steps of (L2+1)/2 makes that partner the unit loaded one step earlier, so only two
    // units are live at a time.
    constexpr int S = (L2 + 1) / 2;
    float4 prev = line[((L2 - 1) * S) % L2];
    static_for<0, L2>([&](auto ic) {
        constexpr int i = decltype(ic)::value;
        constexpr int n2 = (i * S) % L2;
        const float4 cur = line[n2];
        y[n2] = cfma_s(sg, float2{prev.z, prev.w}, float2{cur.x, cur.y});
        prev = cur;
    });
}

template <int RA, int RB>
__device__ __forceinline__ void line_fft(float2 (&y)[RA * RB])
{
    if constexpr (RB == 1) dft_reg<RA>(y);
    else dft_pfa<RA, RB>(y);
}

// real-FFT split of the thread's parity p and transposed store: X[k] -> dst[k*H].  One code path for both
// parities (the kernel is instruction-cache hungry): the twiddles come from a table, and k = ka(p, k2) is a
// compile-time offset from one of two base pointers, dst +- p*Mh*H.
template <class C>
__device__ __forceinline__ void split_store(const float2 (&z)[C::Mh], float2* __restrict__ dst, int p, const float2* __restrict__ tw)
{
    constexpr int M = C::M, Mh = C::Mh, H = C::H;
    if (p == 0) {
        dst[0] = float2{z[0].x + z[0].y, 0.f};
        dst[M * H] = float2{z[0].x - z[0].y, 0.f};
    } else {
        dst[Mh * H] = float2{z[0].x, -z[0].y};
    }
    float2* dp = dst + p * (Mh * H);
    float2* dm = dst - p * (Mh * H);
    static_for<1, C::NPAIR + 1>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        constexpr int c0 = ((Mh + 1) * k2) % M;          // ka = c0 + Mh*p (c0 < Mh) or c0 - Mh*p; kb = M - ka
        const float2 t = tw[k2];                         // (-i/2) * w_W^ka
        const float2 a = z[k2], b = z[Mh - k2];
        const float sx = a.x + b.x, sy = a.y - b.y;      // a + conj b
        const float dx = a.x - b.x, dy = a.y + b.y;      // a - conj b
        const float ux = t.x * dx - t.y * dy, uy = t.x * dy + t.y * dx;
        const float2 xa = float2{fmaf(0.5f, sx, ux), fmaf(0.5f, sy, uy)};
        const float2 xb = float2{fmaf(0.5f, sx, -ux), fmaf(-0.5f, sy, uy)};
        if constexpr (c0 < Mh) {
            dp[c0 * H] = xa;
            dm[(M - c0) * H] = xb;
        } else {
            dm[c0 * H] = xa;
            dp[(M - c0) * H] = xb;
        }
    });
}

// (-i/2) * w_W^ka for the forward split, ka = (Mh*p + (Mh+1)*k2) mod M
template <class C>
struct SplitTable {
    float x[2 * (C::NPAIR + 1)], y[2 * (C::NPAIR + 1)];
    constexpr SplitTable() : x{}, y{}
    {
        for (int p = 0; p < 2; ++p)
            for (int k2 = 0; k2 <= C::NPAIR; ++k2) {
                const int ka = (C::Mh * p + (C::Mh + 1) * k2) % C::M;
                x[p * (C::NPAIR + 1) + k2] = float(-0.5 * cx_sin2pi(ka, C::W));
                y[p * (C::NPAIR + 1) + k2] = float(-0.5 * cx_cos2pi(ka, C::W));
            }
    }
};
template <class C>
__device__ const SplitTable<C> g_split_table{};

// column results of parity p: Y[k2] is output row (Hh*p + (Hh+1)*k2) mod H
template <class C>
__device__ __forceinline__ void store_cols(const float2 (&y)[C::Hh], float2* __restrict__ dst, int p)
{
    float2* dp = dst + p * (C::Hh * C::MO);
    float2* dm = dst - p * (C::Hh * C::MO);
    static_for<0, C::Hh>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        constexpr int c0 = ((C::Hh + 1) * k2) % C::H;
        if constexpr (c0 < C::Hh) dp[c0 * C::MO] = y[k2];
        else dm[c0 * C::MO] = y[k2];
    });
}

// inverse column fold: the forward butterflies run on the index-reversed line x'[n] = x[(L - n) mod L], which
// makes them the inverse transform.  y[n2] = x'[2*n2] + sg * x'[(L2 + 2*n2) mod L], elements STRIDE apart.
template <int L2, int STRIDE>
__device__ __forceinline__ void load_fold_rev(float2 (&y)[L2], const float2* __restrict__ src, float sg)
{
    constexpr int L = 2 * L2;
    static_for<0, L2>([&](auto nc) {
        constexpr int n2 = decltype(nc)::value;
        constexpr int ia = (L - 2 * n2) % L, ib = (2 * L - L2 - 2 * n2) % L;
        const float2 a = src[ia * STRIDE], b = src[ib * STRIDE];
        y[n2] = cfma_s(sg, b, a);
    });
}

// column results of parity P, transposed: value k2 is row r = (Hh*P + (Hh+1)*k2) mod H of this column
template <class C, int P>
__device__ __forceinline__ void store_rows_t(const float2 (&y)[C::Hh], float2* __restrict__ dst)
{
    static_for<0, C::Hh>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        constexpr int r = (C::Hh * P + (C::Hh + 1) * k2) % C::H;
        dst[r] = y[k2];
    });
}

// i * w_W^(-ka) for the Hermitian merge, ka = Mh*p + 2*k2 (compile-time table, scaled at kernel start)
template <class C>
struct MergeTable {
    float x[2 * (C::NPAIR + 1)], y[2 * (C::NPAIR + 1)];
    constexpr MergeTable() : x{}, y{}
    {
        for (int p = 0; p < 2; ++p)
            for (int k2 = 0; k2 <= C::NPAIR; ++k2) {
                const int ka = C::Mh * p + 2 * k2;
                x[p * (C::NPAIR + 1) + k2] = float(-cx_sin2pi(ka, C::W));
                y[p * (C::NPAIR + 1) + k2] = float(cx_cos2pi(ka, C::W));
            }
    }
};
template <class C>
__device__ const MergeTable<C> g_merge_table{};

// ---- truncated modes (spectral layers keep a corner of low modes) --------------------------------------------------
// FNO / AFNO style layers use only the modes rows q in [0, kh) u [H - kh, H), columns k in [0, kw) of the one-sided
// spectrum.  The truncated forward never writes the discarded modes and the truncated inverse never reads them: the
// kept block [2*kh][kw] travels as ONE dense bulk copy (16*kh*kw bytes per image instead of H*MO*8), so the complex
// side of the traffic shrinks by H*MO / (2*kh*kw).  Arithmetic as in the full kernels; columns k >= kw are skipped.

// kept-row index of spectrum row q, or -1
__device__ __forceinline__ int kept_row(int q, int H, int kh) { return q < kh ? q : (q >= H - kh ? q - (H - 2 * kh) : -1); }



// I/O warp of group g (one thread): for the group's images in order, wait until the result is staged, store
// it, and refill the stage.  The store leaves in KP pieces (one bulk group each) and piece j of the next
// image is requested as soon as pieces 0..j have been read out, so the stage's drain and its refill overlap
// instead of adding up -- a stage in transit is what a group waits for.  done[g][s] is private to the
// (group, stage) pair, whose uses are LCM images apart, so its phase parity is (it / LCM) & 1.
// LOAD_BYTES / STORE_BYTES are compile-time constants in the full kernels (everything below folds) and run-time values in
// the truncated-mode kernels, where one side is the small kept block.
template <class C, int STAGE>
__device__ __forceinline__ void io_loop(int g, const unsigned char* __restrict__ in, unsigned char* __restrict__ out, long long nimg,
                                        unsigned char* stages, unsigned full0, unsigned done0, const int LOAD_BYTES, const int STORE_BYTES,
                                        const int LOAD_AT = 0 /* byte offset of the loaded block inside its stage */)
{
    constexpr int NG = C::NG, NS = C::NS, KP = 4;
    const int SP = (STORE_BYTES / KP + 15) / 16 * 16;          // store piece (the last one is shorter, possibly empty)
    const int LP0 = (LOAD_BYTES / KP) / 16 * 16;
    const int LP = LP0 < SP ? LP0 : SP;                        // load piece  (the last one is longer): piece j lies inside store pieces 0..j
    auto load_piece = [&](long long tile, int s, int j) {
        const unsigned off = (unsigned)(j * LP);
        const unsigned bytes = (j == KP - 1) ? (unsigned)(LOAD_BYTES - (KP - 1) * LP) : (unsigned)LP;
        if (bytes == 0) return;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         pipe::s32(stages + (size_t)s * STAGE) + (unsigned)LOAD_AT + off),
                     "l"(in + tile * (long long)LOAD_BYTES + off), "r"(bytes), "r"(full0 + 8 * s)
                     : "memory");
    };
    auto expect = [&](int s) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full0 + 8 * s), "r"((unsigned)LOAD_BYTES) : "memory");
    };
    if (g == 0)
        for (int s = 0; s < NS; ++s) {
            const long long tile = blockIdx.x + s * (long long)gridDim.x;
            if (tile >= nimg) break;
            expect(s);
#pragma unroll
            for (int j = 0; j < KP; ++j) load_piece(tile, s, j);
        }
    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        pipe::mbar_wait(done0 + 8 * (g * NS + s), (unsigned)((it / C::LCM) & 1));
#pragma unroll
        for (int j = 0; j < KP; ++j) {
            const int off = j * SP < STORE_BYTES ? j * SP : STORE_BYTES;
            const int end = (j == KP - 1 || (j + 1) * SP > STORE_BYTES) ? STORE_BYTES : (j + 1) * SP;
            if (end > off)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + tile * (long long)STORE_BYTES + off),
                             "r"(pipe::s32(stages + (size_t)s * STAGE) + (unsigned)off), "r"((unsigned)(end - off))
                             : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");       // (an empty group completes at once)
        }
        const long long next = tile + (long long)NS * gridDim.x;   // the image that uses this stage next
        if (next < nimg) {
            expect(s);
            asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
            load_piece(next, s, 0);
            asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
            load_piece(next, s, 1);
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            load_piece(next, s, 2);
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            load_piece(next, s, 3);
        }
    }
    // the last bulk stores must have completed before the CTA (and its shared memory) goes away
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- 16-bit tensors at the kernel's boundary (trt_dft_plan_create_typed) ---------------------------------------------
// IO = float is the ordinary kernel.  IO = __half / __nv_bfloat16: both tensors hold 16-bit elements, everything between the
// first shared-memory read and the last shared-memory write is the fp32 kernel unchanged.  The real image is H*W*2 bytes,
// a multiple of 16, and travels as before.  The spectrum is H*MO*4 bytes = 8 mod 16 per image: image t starts 8 bytes off a
// 16-byte boundary when t is odd.  Bulk copies need 16-byte alignment at both ends, so an odd image sits 8 bytes into its
// stage, and every image travels as a 16-byte aligned CORE of H*MO*4 - 8 bytes (bulk copy) plus an 8-byte EDGE -- its
// first two elements when t is odd, its last two when t is even -- which the I/O thread moves with one ordinary
// 64-bit access.
using dft::Io16;

// load_fold over a line of 2*L2 complex values whose 8-byte unit j holds x[2j], x[2j+1] as four 16-bit numbers
template <class IO, int L2>
__device__ __forceinline__ void load_fold16(float2 (&y)[L2], const uint2* __restrict__ line, unsigned short sg)
{
    constexpr int S = (L2 + 1) / 2;
    uint2 prev = line[((L2 - 1) * S) % L2];
    static_for<0, L2>([&](auto ic) {
        constexpr int i = decltype(ic)::value;
        constexpr int n2 = (i * S) % L2;
        const uint2 cur = line[n2];
        y[n2] = Io16<IO>::fold(cur.x, prev.y, sg);
        prev = cur;
    });
}

template <class IO, int L2, int STRIDE>
__device__ __forceinline__ void load_fold_rev16(float2 (&y)[L2], const unsigned* __restrict__ src, unsigned short sg)
{
    constexpr int L = 2 * L2;
    static_for<0, L2>([&](auto nc) {
        constexpr int n2 = decltype(nc)::value;
        constexpr int ia = (L - 2 * n2) % L, ib = (2 * L - L2 - 2 * n2) % L;
        y[n2] = Io16<IO>::fold(src[ia * STRIDE], src[ib * STRIDE], sg);
    });
}

// store_cols with 16-bit elements: dst points at element [0][column] of the (possibly shifted) tile [q][k]
template <class IO, class C>
__device__ __forceinline__ void store_cols16(const float2 (&y)[C::Hh], typename Io16<IO>::v2* __restrict__ dst, int p)
{
    typename Io16<IO>::v2* dp = dst + p * (C::Hh * C::MO);
    typename Io16<IO>::v2* dm = dst - p * (C::Hh * C::MO);
    static_for<0, C::Hh>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        constexpr int c0 = ((C::Hh + 1) * k2) % C::H;
        if constexpr (c0 < C::Hh) dp[c0 * C::MO] = Io16<IO>::down(y[k2]);
        else dm[c0 * C::MO] = Io16<IO>::down(y[k2]);
    });
}

// io_loop for the 16-bit kernels.  FWD: loads real images (RB bytes at the start of the stage), stores spectra (core +
// edge).  !FWD: loads spectra, stores real images.  Pieces as in io_loop; a load piece may start 16 bytes into the stage,
// hence LP <= SP - 16.
template <class C, int STAGE, bool FWD>
__device__ __forceinline__ void io_loop16(int g, const unsigned char* __restrict__ in, unsigned char* __restrict__ out, long long nimg,
                                          unsigned char* stages, unsigned full0, unsigned done0)
{
    constexpr int NG = C::NG, NS = C::NS, KP = 4;
    constexpr int RB = C::TILE_F / 2, CB = C::TILE_I / 2, CORE = CB - 8;
    static_assert(RB % 16 == 0 && CB % 16 == 8, "real images travel whole, spectra as core + edge");
    constexpr int LOADB = FWD ? RB : CORE, STOREB = FWD ? CORE : RB;
    constexpr int SP = (STOREB / KP + 15) / 16 * 16;
    constexpr int LP0 = (LOADB / KP) / 16 * 16;
    constexpr int LP = LP0 < SP - 16 ? LP0 : SP - 16;
    static_assert(LP > 0 && (KP - 1) * SP < STOREB && (KP - 1) * LP < LOADB, "piece sizes");
    auto load_piece = [&](long long tile, int s, int j) {
        const unsigned odd = FWD ? 0u : (unsigned)(tile & 1);
        const unsigned off = (unsigned)(j * LP);
        const unsigned bytes = (j == KP - 1) ? (unsigned)(LOADB - (KP - 1) * LP) : (unsigned)LP;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         pipe::s32(stages + (size_t)s * STAGE) + 16u * odd + off),
                     "l"(in + tile * (long long)(FWD ? RB : CB) + 8u * odd + off), "r"(bytes), "r"(full0 + 8 * s)
                     : "memory");
    };
    auto expect = [&](int s) {
        asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(full0 + 8 * s), "r"((unsigned)LOADB) : "memory");
    };
    // spectrum edge of image `tile`: global byte offset inside the image / shared byte offset inside the stage
    auto edge_g = [&](long long tile) { return (tile & 1) ? 0 : CORE; };
    auto edge_s = [&](long long tile) { return (tile & 1) ? 8 : CORE; };
    auto load_edge = [&](long long tile) {
        return __ldg(reinterpret_cast<const uint2*>(in + tile * (long long)CB + edge_g(tile)));
    };
    auto put_edge_and_arrive = [&](long long tile, int s, uint2 e) {
        if (!FWD) *reinterpret_cast<uint2*>(stages + (size_t)s * STAGE + edge_s(tile)) = e;
        pipe::mbar_arrive(full0 + 8 * s);          // (release: the edge is visible to whoever sees the phase complete)
    };
    if (g == 0)
        for (int s = 0; s < NS; ++s) {
            const long long tile = blockIdx.x + s * (long long)gridDim.x;
            if (tile >= nimg) break;
            uint2 e = uint2{0u, 0u};
            if (!FWD) e = load_edge(tile);
            expect(s);
#pragma unroll
            for (int j = 0; j < KP; ++j) load_piece(tile, s, j);
            put_edge_and_arrive(tile, s, e);
        }
    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        const long long next = tile + (long long)NS * gridDim.x;   // the image that uses this stage next
        uint2 e = uint2{0u, 0u};
        if (!FWD && next < nimg) e = load_edge(next);              // in flight while this image is still being computed
        pipe::mbar_wait(done0 + 8 * (g * NS + s), (unsigned)((it / C::LCM) & 1));
        const unsigned odd = FWD ? (unsigned)(tile & 1) : 0u;
        unsigned char* stage = stages + (size_t)s * STAGE;
        if (FWD) *reinterpret_cast<uint2*>(out + tile * (long long)CB + edge_g(tile)) = *reinterpret_cast<const uint2*>(stage + edge_s(tile));
#pragma unroll
        for (int j = 0; j < KP; ++j) {
            const int off = j * SP;
            const int end = (j == KP - 1) ? STOREB : (j + 1) * SP;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + tile * (long long)(FWD ? CB : RB) + 8u * odd + off),
                         "r"(pipe::s32(stage) + 16u * odd + (unsigned)off), "r"((unsigned)(end - off))
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        if (next < nimg) {
            expect(s);
            asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
            load_piece(next, s, 0);
            asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
            load_piece(next, s, 1);
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            load_piece(next, s, 2);
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            load_piece(next, s, 3);
            put_edge_and_arrive(next, s, e);       // (every store piece has been read out: the edge may land anywhere)
        }
    }
    // the last bulk stores must have completed before the CTA (and its shared memory) goes away
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace reg2d

// ---- forward ---------------------------------------------------------------------------------
// TR: truncated modes (see above io_loop): only rows [0, kh) u [H - kh, H), columns [0, kw) of the spectrum leave
// the kernel, as a dense [2*kh][kw] block; columns past kw are not transformed at all.
template <class C, bool TR, class IO = float>
__device__ __forceinline__ void reg_rfft2_body(const float* __restrict__ in, float* __restrict__ out, long long nimg, const int kh, const int kw)
{
    constexpr bool IO16 = !std::is_same<IO, float>::value;
    static_assert(!TR || C::IOW, "the truncated kernels use the I/O-warp layout");
    static_assert(!IO16 || (C::IOW && !TR), "the 16-bit kernels use the I/O-warp layout and move the full spectrum");
    // (truncated: the two halves of an image's work are no longer equal, so the groups run free instead of taking turns)
    constexpr int PPX = TR ? 0 : C::PP;
    constexpr int H = C::H, M = C::M, MO = C::MO, GT = C::GT, NG = C::NG, NS = C::NS, WPP = C::WPP;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* stages = smem_raw;
    float2* tws = reinterpret_cast<float2*>(smem_raw + NS * C::STAGE_F);      // (-i/2) * w_W^ka, [parity][k2]
    const unsigned full0 = pipe::s32(tws + C::NTAB_F);              // image has landed in the stage
    const unsigned done0 = full0 + 8 * NS;                          // [group][stage] spectrum staged (IOW only)
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB_F; i += C::THREADS) tws[i] = float2{reg2d::g_split_table<C>.x[i], reg2d::g_split_table<C>.y[i]};
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) pipe::mbar_init(full0 + 8 * s, 1);
        for (int s = 0; s < NG * NS; ++s) pipe::mbar_init(done0 + 8 * s, GT / 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pipe::fence_async();
    }
    __syncthreads();

    if constexpr (C::IOW) {
        if (tid >= NG * GT) {
            if ((tid & 31) == 0) {
                if constexpr (IO16)
                    reg2d::io_loop16<C, C::STAGE_F, true>((tid - NG * GT) >> 5, reinterpret_cast<const unsigned char*>(in),
                                                          reinterpret_cast<unsigned char*>(out), nimg, stages, full0, done0);
                else
                    reg2d::io_loop<C, C::STAGE_F>((tid - NG * GT) >> 5, reinterpret_cast<const unsigned char*>(in),
                                                  reinterpret_cast<unsigned char*>(out), nimg, stages, full0, done0, C::TILE_F,
                                                  TR ? 16 * kh * kw : C::TILE_I);
            }
            return;
        }
    }
    // No I/O warp (12 warps keep 168 registers each).  Thread 0 fills the ring once; afterwards the first
    // thread of each group stores the group's finished spectrum with one bulk copy and, one phase later,
    // refills that stage with the image that uses it next (image it + NS).
    auto issue_load = [&](long long it_load, int s) {
        const long long tile = blockIdx.x + it_load * (long long)gridDim.x;
        if (tile >= nimg) return;
        const unsigned bar = full0 + 8 * s;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)C::TILE_F) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         pipe::s32(stages + (size_t)s * C::STAGE_F)),
                     "l"(reinterpret_cast<const unsigned char*>(in) + tile * (long long)C::TILE_F), "r"((unsigned)C::TILE_F), "r"(bar)
                     : "memory");
    };
    if (!C::IOW && tid == 0)
        for (int s = 0; s < NS; ++s) issue_load(s, s);
    const int g = tid / GT;
    const int gt = tid - g * GT;
    const int warp = gt >> 5;
    const int p = warp / WPP;                           // parity handled by this warp
    const int line = (warp - p * WPP) * 32 + (gt & 31); // row in the row phase, column in the column phase
    const float sg = p ? -1.f : 1.f;
    // lanes past the last line redo the last line (same values to the same addresses): no divergence, and
    // no conditionally-live register arrays across the barriers
    const int lrow = line < H ? line : H - 1, lcol = line < MO ? line : MO - 1;
    const int bar_id = 1 + g;
    long long pend_it = -1;                             // gt == 0: image whose bulk store is still reading its stage
    auto refill = [&]() {
        if (!C::IOW && gt == 0 && pend_it >= 0) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            issue_load(pend_it + NS, (int)(pend_it % NS));
            pend_it = -1;
        }
    };

    if (PPX != 0 && g == 1) asm volatile("bar.arrive %0, %1;" ::"r"(3), "r"(2 * GT) : "memory");   // group 0 goes first
    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        float2* buf = reinterpret_cast<float2*>(stages + (size_t)s * C::STAGE_F);
        DFT_TR(it, 0);
        if (PPX == 1) pipe::group_sync(3 + g, 2 * GT);     // my turn
        pipe::mbar_wait(full0 + 8 * s, (unsigned)((it / NS) & 1));
        DFT_TR(it, 1);

        // rows: the whole packed row -> registers; the transposed result [k][row] overwrites the stage
        {
            float2 z[C::Mh];
            if constexpr (IO16)
                reg2d::load_fold16<IO, C::Mh>(z, reinterpret_cast<const uint2*>(buf) + lrow * (M / 2),
                                              p ? reg2d::Io16<IO>::kMinusOne : reg2d::Io16<IO>::kOne);
            else reg2d::load_fold<C::Mh>(z, reinterpret_cast<const float4*>(buf) + lrow * (M / 2), sg);
            refill();                                   // the previous image's store has long since drained
            DFT_TR(it, 2);
            pipe::group_sync(bar_id, GT);               // every row has been read
            DFT_TR(it, 3);
            reg2d::line_fft<C::RA, C::RB>(z);
            reg2d::split_store<C>(z, buf + lrow, p, tws + p * (C::NPAIR + 1));
        }
        DFT_TR(it, 4);
        if (PPX == 1) asm volatile("bar.arrive %0, %1;" ::"r"(4 - g), "r"(2 * GT) : "memory");   // the other group's turn
        pipe::group_sync(bar_id, GT);
        DFT_TR(it, 5);
        // columns: column k is a contiguous line of the transposed tile; the spectrum [q][k] overwrites it
        // (truncated: a warp none of whose columns is kept skips the column transform)
        if (!TR || (warp - p * WPP) * 32 < kw) {
            float2 y[C::Hh];
            reg2d::load_fold<C::Hh>(y, reinterpret_cast<const float4*>(buf) + lcol * (H / 2), sg);
            DFT_TR(it, 6);
            reg2d::line_fft<C::CA, C::CB>(y);
            pipe::group_sync(bar_id, GT);               // every column has been read
            if constexpr (IO16) {
                typedef typename reg2d::Io16<IO>::v2 v2;
                reg2d::store_cols16<IO, C>(y, reinterpret_cast<v2*>(reinterpret_cast<unsigned char*>(buf) + 8 * (int)(tile & 1)) + lcol, p);
            } else {
                reg2d::store_cols<C>(y, buf + lcol, p);
            }
        } else {
            pipe::group_sync(bar_id, GT);
        }
        {
            if constexpr (TR) {
                // Compact the kept modes, rows [0, kh) u [H - kh, H) x columns [0, kw) of the tile [q][k] (pitch MO), into
                // a dense block at the start of the stage, GT elements at a time: element e comes from an index >= e, so
                // reading a chunk, a barrier, and writing it never overwrites anything that is still to be read.
                pipe::group_sync(bar_id, GT);           // the spectrum tile is complete
                const int total = 2 * kh * kw;
                const unsigned kwm = dft::magic_of((unsigned)kw);
                constexpr int EPT = 4;                  // elements per thread and chunk
                for (int base = 0; base < total; base += EPT * GT) {
                    float2 v[EPT];
#pragma unroll
                    for (int j = 0; j < EPT; ++j) {
                        const int e = base + j * GT + gt;
                        const int ec = e < total ? e : total - 1;
                        const int qi = dft::fdiv(ec, kwm), k = ec - qi * kw;
                        v[j] = buf[(qi < kh ? qi : qi + (H - 2 * kh)) * MO + k];
                    }
                    pipe::group_sync(bar_id, GT);
#pragma unroll
                    for (int j = 0; j < EPT; ++j) {
                        const int e = base + j * GT + gt;
                        if (e < total) buf[e] = v[j];
                    }
                }
            }
            pipe::fence_async();
            if constexpr (C::IOW) {
                __syncwarp();
                if ((gt & 31) == 0) pipe::mbar_arrive(done0 + 8 * (g * NS + s));
                DFT_TR(it, 7);
                continue;
            }
            pipe::group_sync(bar_id, GT);
            if (gt == 0) {
                float* dst = out + tile * (long long)(H * MO * 2);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(pipe::s32(buf)),
                             "r"((unsigned)C::TILE_I)
                             : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                pend_it = it;
            }
            DFT_TR(it, 7);
        }
    }
    refill();
    // the last bulk store must have completed before the CTA (and its shared memory) goes away
    if (!C::IOW && gt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_rfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                     const float2* __restrict__ /*tables*/, float /*scale*/)
{
    reg_rfft2_body<C, false>(in, out, nimg, 0, 0);
}
template <class C, class IO>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_rfft2_io16_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg, const float2* __restrict__ /*tables*/,
                          float /*scale*/)
{
    reg_rfft2_body<C, false, IO>(in, out, nimg, 0, 0);
}
template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_rfft2_trunc_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg, const float2* __restrict__ /*tables*/,
                           float /*scale*/, int kh, int kw)
{
    reg_rfft2_body<C, true>(in, out, nimg, kh, kw);
}

// ---- inverse ---------------------------------------------------------------------------------
template <class C, bool TR, class IO = float>
__device__ __forceinline__ void reg_irfft2_body(const float* __restrict__ in, float* __restrict__ out, long long nimg, float scale,
                                                const int kh, const int kw)
{
    constexpr bool IO16 = !std::is_same<IO, float>::value;
    static_assert(!TR || C::IOW, "the truncated kernels use the I/O-warp layout");
    static_assert(!IO16 || (C::IOW && !TR), "the 16-bit kernels use the I/O-warp layout and move the full spectrum");
    constexpr int PPX = TR ? 0 : C::PP;
    constexpr int H = C::H, M = C::M, Mh = C::Mh, MO = C::MO, PI = C::PI, GT = C::GT, NG = C::NG, NS = C::NS, WPP = C::WPP;
    constexpr int NTW = C::NPAIR + 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char* stages = smem_raw;
    float2* twm = reinterpret_cast<float2*>(smem_raw + NS * C::STAGE_I);      // scale * i * w_W^(-ka), [parity][k2]
    const unsigned full0 = pipe::s32(twm + C::NTAB_I);
    const unsigned done0 = full0 + 8 * NS;                          // [group][stage] image staged (IOW only)
    const int tid = threadIdx.x;

    for (int i = tid; i < C::NTAB_I; i += C::THREADS)
        twm[i] = float2{scale * reg2d::g_merge_table<C>.x[i], scale * reg2d::g_merge_table<C>.y[i]};
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) pipe::mbar_init(full0 + 8 * s, 1);
        for (int s = 0; s < NG * NS; ++s) pipe::mbar_init(done0 + 8 * s, GT / 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        pipe::fence_async();
    }
    __syncthreads();

    if constexpr (C::IOW) {
        if (tid >= NG * GT) {
            if ((tid & 31) == 0) {
                if constexpr (IO16)
                    reg2d::io_loop16<C, C::STAGE_I, false>((tid - NG * GT) >> 5, reinterpret_cast<const unsigned char*>(in),
                                                           reinterpret_cast<unsigned char*>(out), nimg, stages, full0, done0);
                else
                    reg2d::io_loop<C, C::STAGE_I>((tid - NG * GT) >> 5, reinterpret_cast<const unsigned char*>(in),
                                                  reinterpret_cast<unsigned char*>(out), nimg, stages, full0, done0,
                                                  TR ? 16 * kh * kw : C::TILE_I, C::TILE_F, TR ? (C::TAIL - 2 * kh * kw) * 8 : 0);
            }
            return;
        }
    }
    auto issue_load = [&](long long it_load, int s) {
        const long long tile = blockIdx.x + it_load * (long long)gridDim.x;
        if (tile >= nimg) return;
        const unsigned bar = full0 + 8 * s;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)C::TILE_I) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         pipe::s32(stages + (size_t)s * C::STAGE_I)),
                     "l"(reinterpret_cast<const unsigned char*>(in) + tile * (long long)C::TILE_I), "r"((unsigned)C::TILE_I), "r"(bar)
                     : "memory");
    };
    if (!C::IOW && tid == 0)
        for (int s = 0; s < NS; ++s) issue_load(s, s);
    const int g = tid / GT;
    const int gt = tid - g * GT;
    const int warp = gt >> 5, lane = gt & 31;
    // column phase: parity per warp, lanes along the columns
    const int pc = warp / WPP;
    const int col = (warp - pc * WPP) * 32 + lane;
    const int lcol = col < MO ? col : MO - 1;
    const float sgc = pc ? -1.f : 1.f;
    // row phase: the two parities of a row sit 16 lanes apart, so the last radix-2 step is one shuffle
    const int pr = lane >> 4;
    const int row = warp * 16 + (lane & 15);
    const int lrow = row < H ? row : H - 1;
    const float sgr = pr ? -1.f : 1.f;
    const int bar_id = 1 + g;
    long long pend_it = -1;                             // gt == 0: image whose bulk store is still reading its stage
    auto refill = [&]() {
        if (!C::IOW && gt == 0 && pend_it >= 0) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            issue_load(pend_it + NS, (int)(pend_it % NS));
            pend_it = -1;
        }
    };

    if (PPX != 0 && g == 1) asm volatile("bar.arrive %0, %1;" ::"r"(3), "r"(2 * GT) : "memory");   // group 0 goes first
    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < nimg; tile += (long long)NG * gridDim.x, it += NG) {
        const int s = (int)(it % NS);
        float2* buf = reinterpret_cast<float2*>(stages + (size_t)s * C::STAGE_I);
        if (PPX == 1) pipe::group_sync(3 + g, 2 * GT);     // my turn
        pipe::mbar_wait(full0 + 8 * s, (unsigned)((it / NS) & 1));

        // columns: column k of the spectrum [q][k] -> registers; the transposed result [k][row], pitch PI,
        // overwrites the stage
        if constexpr (TR) {
            // The kept block [2*kh][kw] sits at elements [TAIL - total, TAIL) of the stage.  Expand it in place into the
            // spectrum tile [q][k] (pitch MO), zeros everywhere else: an element's place in the tile never lies behind
            // its place in the block, so chunks go front to back (read, barrier, write).
            const int total = 2 * kh * kw, blk0 = C::TAIL - total;
            const float4 z4 = float4{0.f, 0.f, 0.f, 0.f};
            for (int i = gt; i < blk0 / 2; i += GT) reinterpret_cast<float4*>(buf)[i] = z4;      // ahead of the block: all zero for now
            pipe::group_sync(bar_id, GT);
            const unsigned kwm = dft::magic_of((unsigned)kw);
            constexpr int EPT = 4;
            for (int base = 0; base < total; base += EPT * GT) {
                float2 v[EPT];
#pragma unroll
                for (int j = 0; j < EPT; ++j) {
                    const int e = base + j * GT + gt;
                    v[j] = buf[blk0 + (e < total ? e : total - 1)];
                }
                pipe::group_sync(bar_id, GT);
#pragma unroll
                for (int j = 0; j < EPT; ++j) {
                    const int e = base + j * GT + gt;
                    if (e < total) {
                        const int qi = dft::fdiv(e, kwm), k = e - qi * kw;
                        buf[(qi < kh ? qi : qi + (H - 2 * kh)) * MO + k] = v[j];
                    }
                }
            }
            pipe::group_sync(bar_id, GT);
            // what is left of the block's own footprint and is not a kept mode goes back to zero
            constexpr unsigned mom = dft::magic_u32((unsigned)MO);
            for (int pos = blk0 + gt; pos < H * MO; pos += GT) {
                const int q = dft::fdiv(pos, mom), k = pos - q * MO;
                if (!(k < kw && (q < kh || q >= H - kh))) buf[pos] = float2{0.f, 0.f};
            }
            pipe::group_sync(bar_id, GT);
        }
        if (TR && (warp - pc * WPP) * 32 >= kw) {
            // truncated: every column of this warp is zero, and so is its transform
            refill();
            pipe::group_sync(bar_id, GT);
            float2* dst = buf + lcol * PI;
            static_for<0, C::Hh>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                constexpr int r0 = ((C::Hh + 1) * k2) % H;                      // parity 0 writes row r0, parity 1 row (r0 + Hh) mod H
                dst[pc ? (r0 + C::Hh) % H : r0] = float2{0.f, 0.f};
            });
        } else {
            float2 y[C::Hh];
            if constexpr (IO16) {
                reg2d::load_fold_rev16<IO, C::Hh, MO>(
                    y, reinterpret_cast<const unsigned*>(reinterpret_cast<const unsigned char*>(buf) + 8 * (int)(tile & 1)) + lcol,
                    pc ? reg2d::Io16<IO>::kMinusOne : reg2d::Io16<IO>::kOne);
            } else {
                reg2d::load_fold_rev<C::Hh, MO>(y, buf + lcol, sgc);
            }
            refill();                                   // the previous image's store has long since drained
            pipe::group_sync(bar_id, GT);               // every column has been read
            reg2d::line_fft<C::CA, C::CB>(y);
            if (pc == 0) reg2d::store_rows_t<C, 0>(y, buf + lcol * PI);
            else reg2d::store_rows_t<C, 1>(y, buf + lcol * PI);
        }
        if (PPX == 1) asm volatile("bar.arrive %0, %1;" ::"r"(4 - g), "r"(2 * GT) : "memory");   // the other group's turn
        pipe::group_sync(bar_id, GT);
        // rows: Hermitian merge of this thread's parity -> FFT -> radix-2 step across the half warps
        {
            float2 z[Mh];
            const float2* pa = buf + lrow + pr * (Mh * PI);     // X[ka], ka = Mh*pr + 2*k2
            const float2* pb = buf + lrow - pr * (Mh * PI);     // X[kb], kb = M - ka
            const float2* tw = twm + pr * NTW;
            {
                // k2 = 0: Z[0] from the (real) DC and Nyquist bins, or Z[Mh] = 2 conj X[Mh]
                const float2 u = pa[0], v = pb[M * PI];
                z[0] = float2{scale * (u.x + v.x), scale * (pr ? -(u.y + v.y) : (u.x - v.x))};
            }
            static_for<1, NTW>([&](auto kc) {
                constexpr int k2 = decltype(kc)::value;
                const float2 a = pa[2 * k2 * PI], b = pb[(M - 2 * k2) * PI];
                const float2 t = tw[k2];
                const float sx = a.x + b.x, sy = a.y - b.y;      // a + conj b
                const float dx = a.x - b.x, dy = a.y + b.y;      // a - conj b
                const float ux = t.x * dx - t.y * dy, uy = t.x * dy + t.y * dx;
                // Z[ka] = S + T goes to slot k2, Z[kb] = conj(S - T) to slot Mh - k2; the slots are filled
                // index-reversed so that the forward butterflies compute the inverse transform
                z[Mh - k2] = float2{fmaf(scale, sx, ux), fmaf(scale, sy, uy)};
                z[k2] = float2{fmaf(scale, sx, -ux), fmaf(-scale, sy, uy)};
            });
            pipe::group_sync(bar_id, GT);               // every row has been read
            reg2d::line_fft<C::RA, C::RB>(z);
            // z[n2] now holds A_pr[n2]; x = A_0 + (-1)^n1 A_1 at n = (Mh*n1 + (Mh+1)*n2) mod M, n1 = pr
            float2* op = buf + lrow * M + pr * Mh;
            float2* om = buf + lrow * M - pr * Mh;
            static_for<0, Mh>([&](auto nc) {
                constexpr int n2 = decltype(nc)::value;
                constexpr int c0 = ((Mh + 1) * n2) % M;
                const float ox = __shfl_xor_sync(0xffffffffu, z[n2].x, 16);
                const float oy = __shfl_xor_sync(0xffffffffu, z[n2].y, 16);
                const float2 r = float2{fmaf(sgr, z[n2].x, ox), fmaf(sgr, z[n2].y, oy)};
                if constexpr (IO16) {
                    // (the same element of the dense image, four bytes wide instead of eight)
                    typedef typename reg2d::Io16<IO>::v2 v2;
                    if constexpr (c0 < Mh) reinterpret_cast<v2*>(buf)[(op - buf) + c0] = reg2d::Io16<IO>::down(r);
                    else reinterpret_cast<v2*>(buf)[(om - buf) + c0] = reg2d::Io16<IO>::down(r);
                } else {
                    if constexpr (c0 < Mh) op[c0] = r;
                    else om[c0] = r;
                }
            });
            pipe::fence_async();
            if constexpr (C::IOW) {
                __syncwarp();
                if ((gt & 31) == 0) pipe::mbar_arrive(done0 + 8 * (g * NS + s));
                continue;
            }
            pipe::group_sync(bar_id, GT);
            if (gt == 0) {
                float* dst = out + tile * (long long)(H * C::W);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(pipe::s32(buf)),
                             "r"((unsigned)C::TILE_F)
                             : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                pend_it = it;
            }
        }
    }
    refill();
    // the last bulk store must have completed before the CTA (and its shared memory) goes away
    if (!C::IOW && gt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_irfft2_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg,
                      const float2* __restrict__ /*tables*/, float scale)
{
    reg_irfft2_body<C, false>(in, out, nimg, scale, 0, 0);
}
template <class C, class IO>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_irfft2_io16_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg, const float2* __restrict__ /*tables*/,
                           float scale)
{
    reg_irfft2_body<C, false, IO>(in, out, nimg, scale, 0, 0);
}
template <class C>
__global__ void __launch_bounds__(C::THREADS, 1)
    reg_irfft2_trunc_kernel(const float* __restrict__ in, float* __restrict__ out, long long nimg, const float2* __restrict__ /*tables*/,
                            float scale, int kh, int kw)
{
    reg_irfft2_body<C, true>(in, out, nimg, scale, kh, kw);
}

#endif  // __CUDACC__


}  // namespace dft
