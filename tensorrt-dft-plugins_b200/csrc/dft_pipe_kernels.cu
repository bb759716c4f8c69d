// dft_pipe_kernels.cu -- the persistent pipelined shell of the generic engine.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <type_traits>

#include "dft_launch.h"
#include "dft_planner.h"
#include "dft_programs.cuh"
#include "dft_twostep.cuh"

namespace {

using dft::KParams;

__device__ __forceinline__ unsigned smem_u32(const void* p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}

// ---- persistent pipelined variant of the generic engine -----------------------------------------
// One CTA per SM: 2 producer warps stream tiles into a ring of `ns` shared-memory stages (cp.async /
// LDGSTS element copies, or one TMA bulk copy when the tile is contiguous and 16-byte clean), `ng`
// compute groups run the ordinary tile programs (LOAD = false) on their stage and hand it back.
// Used for the row and strided-axis passes of images too large to fuse (BASELINE cfg4 / cfg5).
using dft::kPipeCompute;
using dft::kPipeProducers;
using dft::kPipeThreads;

// (const object with const methods: the programs take their context by reference, and a mutable one whose address has been
// taken is re-read from local memory after every barrier's memory clobber -- a long-scoreboard stall per barrier)
struct GroupCtx {
    const int tid, nthr, bar_id;
    float2* const stg;                  // this group's staging rows for bulk stores (KParams::pack_bulk), else null
    __host__ __device__ __forceinline__ void sync() const
    {
#if defined(__CUDA_ARCH__)
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(nthr) : "memory");
#endif
    }
    __host__ __device__ __forceinline__ bool bulk_load(void*, const void*, long long) const { return false; }
    // bulk stores shared -> global (possibly a peer GPU's memory): issued and awaited by one thread of the group
    __host__ __device__ __forceinline__ float2* stage_out() const { return stg; }
    __host__ __device__ __forceinline__ void fence_async_shared() const
    {
#if defined(__CUDA_ARCH__)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
    }
    __host__ __device__ __forceinline__ void bulk_store(void* dst, const void* src, unsigned bytes) const
    {
#if defined(__CUDA_ARCH__)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
#endif
    }
    __host__ __device__ __forceinline__ void commit_bulk_stores() const
    {
#if defined(__CUDA_ARCH__)
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
#endif
    }
    __host__ __device__ __forceinline__ void wait_bulk_stores_read() const
    {
#if defined(__CUDA_ARCH__)
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
#endif
    }
};

__device__ __forceinline__ void cp_async8(void* dst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void gp_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "GP_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GP_DONE;\n"
        "bra GP_WAIT;\n"
        "GP_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}

// producer side: start the copies of one tile; returns the byte count of a bulk copy issued by
// thread 0 (0 = element copies only)
template <int MODE, int NP>
__device__ __forceinline__ unsigned produce_tile(const KParams& p, long long tile, float2* buf, const float* in, int pt,
                                                 unsigned full_bar, const CUtensorMap* tmap)
{
    if (MODE == dft::MODE_C2C_COLS) {
        const int N = p.H, P = p.P;
        const long long o_ = tile / p.chunks, c_ = tile - o_ * p.chunks, col0 = c_ * p.G;
        if (p.tma_cols) {
            // The whole tile -- N rows of G complex columns out of the padded workspace -- with one tensor copy: the TMA unit
            // walks the strided rows, columns past the end of the workspace row are zero-filled (and dropped later).
            const unsigned bytes = (unsigned)N * (unsigned)p.G * 8u;
            if (pt == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full_bar), "r"(bytes) : "memory");
                asm volatile(
                    "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                        smem_u32(buf)),
                    "l"(tmap), "r"((int)(col0 * 2)), "r"((int)(o_ * N)), "r"(full_bar)
                    : "memory");
            }
            return bytes;
        }
        const int ncol = (int)((p.inner - col0) < p.G ? (p.inner - col0) : p.G);
        const int n1s = p.n1split > 1 ? p.n1split : 1;   // four-step first pass: rows n1 + n1split*n2 of image o_hi
        const long long o_hi = o_ / n1s;
        const int n1 = (int)(o_ - o_hi * n1s);
        const long long ntot = (long long)n1s * N;
        const float2* img = (const float2*)in + o_hi * ntot * p.inner + col0;
        auto src_row = [&](int n) -> const float2* {
            long long row = n1 + (long long)n1s * n;
            if (p.rev_in) row = (row == 0) ? 0 : ntot - row;
            return img + row * p.inner;
        };
        if (NP % ncol == 0) {
            const int rstep = NP / ncol, j0 = pt % ncol, n0 = pt / ncol;
            float2* d0 = buf + n0 * P + j0;
            const int dstep = rstep * P;
            if (!p.rev_in) {
                // constant row stride (n1split * inner): walk pointers, no multiplies in the loop
                const long long rs = (long long)n1s * p.inner;
                const float2* s0 = img + ((long long)n1 + (long long)n1s * n0) * p.inner + j0;
                const long long sstep = (long long)rstep * rs;
#pragma unroll 4
                for (int n = n0; n < N; n += rstep, d0 += dstep, s0 += sstep) cp_async8(d0, s0);
            } else {
#pragma unroll 4
                for (int n = n0; n < N; n += rstep, d0 += dstep) cp_async8(d0, src_row(n) + j0);
            }
        } else {
            for (int idx = pt; idx < N * ncol; idx += NP) {
                const int n = idx / ncol, j = idx - n * ncol;
                cp_async8(buf + n * P + j, src_row(n) + j);
            }
        }
        return 0;
    } else {
        // contiguous tiles.  rows: F1D reads nrows*M packed-complex (= real pairs), I1D nrows*(M+1) complex;
        // fused 2-D: the same per image row, H rows per image, G images per tile (pitch == row length)
        const int M = p.M;
        const bool fwd_side = (MODE == dft::MODE_F1D || MODE == dft::MODE_F2D);
        const bool two_d = (MODE == dft::MODE_F2D || MODE == dft::MODE_I2D);
        const int per_row = fwd_side ? M : M + 1;
        const int P = two_d ? per_row : p.P;
        const long long first = tile * p.G * (two_d ? p.H : 1);
        const long long units_left = (p.nunits - tile * p.G) < p.G ? (p.nunits - tile * p.G) : p.G;
        const int nrows = (int)(units_left * (two_d ? p.H : 1));
        if (MODE == dft::MODE_I1D && p.pack_peer_on) {
            // Pencil irfft2, exchange fused into the LOAD: row R of this rank's band is spread over the ranks' column slabs
            // (KParams::pack_*: slab j holds columns [c0_j, c0_{j+1}) of every row at pitch pack_pitch[j]); the producers
            // fetch each run straight from its owner's memory with element copies (plain loads over NVLink, no multicast).
            for (int r = 0; r < nrows; ++r) {
                const int R = (int)first + r, b = R / p.pack_hn, rl = R - b * p.pack_hn;
                float2* drow = buf + r * P;
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    if (t >= p.pack_n) break;
                    const int c0 = p.pack_c0[t], c1 = p.pack_c0[t + 1];
                    const float2* srow = p.pack_peer[t] + ((long long)b * p.pack_H + p.pack_h0 + rl) * p.pack_pitch[t] - c0;
                    for (int k = c0 + pt; k < c1; k += NP) cp_async8(drow + dft::phys(k, p.pad), srow + k);
                }
            }
            return 0;
        }
        const float* src = in + first * per_row * 2;
        const long long nbytes = (long long)nrows * per_row * 8;
        if (!p.pad && P == per_row && (((uintptr_t)src) & 15) == 0 && (nbytes & 15) == 0 && nbytes < (1 << 20)) {
            if (pt == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full_bar), "r"((unsigned)nbytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_u32(buf)),
                             "l"(src), "r"((unsigned)nbytes), "r"(full_bar)
                             : "memory");
            }
            return (unsigned)nbytes;
        }
        const unsigned mg = dft::magic_of((unsigned)per_row);
        const int total = nrows * per_row;
        if ((((uintptr_t)src) & 7) == 0) {
            const float2* s2 = (const float2*)src;
            for (int idx = pt; idx < total; idx += NP) {
                const int r = dft::fdiv(idx, mg), k = idx - r * per_row;
                cp_async8(buf + r * P + dft::phys(k, p.pad), s2 + idx);
            }
        } else {   // real rows that are only 4-byte aligned
            for (int idx = pt; idx < total; idx += NP) {
                const int r = dft::fdiv(idx, mg), k = idx - r * per_row;
                float* d = (float*)(buf + r * P + dft::phys(k, p.pad));
                cp_async4(d, src + 2 * idx);
                cp_async4(d + 1, src + 2 * idx + 1);
            }
        }
        return 0;
    }
}

// Compile-time radix schedule of the transformed axis.  Injected into the kernel's private copy of
// KParams, it lets the compiler fold the generic tile programs (radix switch, strides, magic divisors,
// skew) down to immediates for registered lengths; GeoNone keeps everything run-time.
struct GeoNone {
    static constexpr bool kStatic = false;
};
template <int N_, int R0, int R1 = 1, int R2 = 1, int R3 = 1>
struct GeoLine {
    static constexpr bool kStatic = true;
    static constexpr int N = N_;
    static constexpr int nsteps = (R0 > 1) + (R1 > 1) + (R2 > 1) + (R3 > 1);
    static_assert(R0 * R1 * R2 * R3 == N_, "schedule must multiply to N");
    static constexpr bool pow2 = (N_ & (N_ - 1)) == 0;
    __device__ static void apply(dft::LineFFT& f)
    {
        constexpr int r[4] = {R0, R1, R2, R3};
        f.n = N_;
        f.nsteps = nsteps;
        int L = N_;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i >= nsteps) break;
            f.radix[i] = r[i];
            const unsigned m = (unsigned)(L / r[i]), bpl = (unsigned)(N_ / r[i]);
            f.mmag[i] = m <= 1 ? 0u : (unsigned)(((1ull << 32) + m - 1) / m);
            f.bmag[i] = bpl <= 1 ? 0u : (unsigned)(((1ull << 32) + bpl - 1) / bpl);
            f.ord[i] = 0;
            L /= r[i];
        }
    }
    // the planner's bank-conflict-free work-item orders for unskewed rows of pitch PITCH, folded at compile time
    template <int PITCH>
    struct Orders {
        dft::LineFFT f;
        constexpr Orders() : f{}
        {
            constexpr int r[4] = {R0, R1, R2, R3};
            f.n = N_;
            f.nsteps = nsteps;
            for (int i = 0; i < nsteps; ++i) f.radix[i] = r[i];
            dft::set_row_orders(f, PITCH);
        }
    };
    template <int PITCH>
    __device__ static void apply_orders(dft::LineFFT& f)
    {
        constexpr Orders<PITCH> o{};
        static_for_orders<PITCH, 0>(f);
    }
    template <int PITCH, int I>
    __device__ static void static_for_orders(dft::LineFFT& f)
    {
        if constexpr (I < nsteps) {
            constexpr Orders<PITCH> o{};
            f.ord[I] = o.f.ord[I];
            f.mmag[I] = o.f.mmag[I];
            f.bmag[I] = o.f.bmag[I];
            static_for_orders<PITCH, I + 1>(f);
        }
    }
};

// Registered strided-axis length with a two-step register-resident program (dft_twostep.cuh): N = (R1A*R1B) * (R2A*R2B),
// tiles TM columns wide.  The generic schedule S0..S2 is still what the planner chose (it keys the dispatch).
template <int N_, int S0, int S1, int S2, int R1A, int R1B, int R2A, int R2B, int TM_, bool FOUR_>
struct GeoTwo : GeoLine<N_, S0, S1, S2, 1> {
    using two_step_cols_tag = void;                        // detected by geo_two_step
    static_assert(R1A * R1B * R2A * R2B == N_, "two-step factors must multiply to N");
    template <class Ctx>
    __device__ static void run_cols(const KParams& p, long long tile, float2* buf, float* out, Ctx& ctx)
    {
        const long long o_ = tile / p.chunks, c_ = tile - o_ * p.chunks, col0 = c_ * p.G;
        const int ncol = (int)((p.inner - col0) < p.G ? (p.inner - col0) : p.G);
        // output rows: pitch out_rowlen (first pass into a padded workspace), or -- second pass out of one -- the image
        // pitch times the rows folded into one workspace row, with the column unfolded per work item (KParams::unflat_in)
        long long orow = p.out_rowlen ? p.out_rowlen : p.inner;
        float2* dst = (float2*)out + o_ * N_ * orow + col0;
        if (p.unflat_in) {
            orow = (p.inner / p.unflat_in) * p.unflat_w;
            dst = (float2*)out + o_ * N_ * orow;
        }
        const int n1s = p.n1split > 1 ? p.n1split : 1;
        const int n1 = (int)(o_ % n1s);
        dft::cols_two_step<R1A, R1B, R2A, R2B, TM_, FOUR_>(buf, p.P, p.tw_col, dst, orow, ncol, p.scale, p.inverse != 0,
                                                    p.n1split > 1 ? p.tw_four : nullptr, n1, ctx,
                                                    p, p.split_n != 0 && p.n1split <= 1, o_, col0);
    }
};
// Registered row length (M packed complex points) with the two-step register-resident row programs.
template <int N_, int S0, int S1, int S2, int R1A, int R1B, int R2A, int R2B>
struct GeoTwoRows : GeoLine<N_, S0, S1, S2, 1> {
    using two_step_rows_tag = void;                        // detected by geo_two_rows
    static constexpr int R1 = R1A * R1B, R2 = R2A * R2B;
    static_assert(R1 * R2 == N_, "two-step factors must multiply to N");
    // T[a*R1 + k1] = w_M^(a*k1), built once per CTA where the generic kernels keep w_M^j
    __device__ static void fill_table(float2* T, const float2* tw_global, int tid, int nthr)
    {
        for (int i = tid; i < N_; i += nthr) {
            const int a = i / R1, k1 = i - a * R1;
            T[i] = __ldg(tw_global + (a * k1) % N_);
        }
    }
    template <class Ctx>
    __device__ static void run_f1d(const KParams& p, long long tile, float2* buf, float* out, Ctx& ctx)
    {
        const long long first = tile * p.G;
        const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
        dft::rows_r2c_two_step<R1A, R1B, R2A, R2B>(buf, nrows, p.P, p.tw_row, p.tw_unt, (float2*)out + first * (N_ + 1), ctx);
    }
    template <class Ctx>
    __device__ static void run_i1d(const KParams& p, long long tile, float2* buf, float* out, Ctx& ctx)
    {
        const long long first = tile * p.G;
        const int nrows = (int)((p.nunits - first) < p.G ? (p.nunits - first) : p.G);
        dft::rows_c2r_two_step<R1A, R1B, R2A, R2B>(buf, nrows, p.P, p.tw_row, p.tw_unt, (float2*)out + first * N_, p.scale, ctx);
    }
};
template <class G, class = void>
struct geo_two_rows : std::false_type {};
template <class G>
struct geo_two_rows<G, std::void_t<typename G::two_step_rows_tag>> : std::true_type {};
template <class G, class = void>
struct geo_two_step : std::false_type {};
template <class G>
struct geo_two_step<G, std::void_t<typename G::two_step_cols_tag>> : std::true_type {};

// CT compute threads + PT producer threads.  The default 512 + 64 = 18 warps puts 5 warps on one of the SM's four
// sub-partitions, which caps a thread at 96 registers (5 x 32 x 96 <= 16 K); 480 + 32 = 16 warps leave 128 and remove the
// 720-point column program's 16 bytes of spills (loop invariants reloaded on the path to every barrier) -- but one producer
// warp cannot keep the tile ring full: cfg4 columns 2.67 -> 2.89 ms.  Kept as an A/B (TRT_DFT_480=1).
template <int MODE, class Geo = GeoNone, int CT = kPipeCompute, int PT = kPipeProducers>
__global__ void __launch_bounds__(CT + PT, 1)
    pipe_tile_kernel(const __grid_constant__ KParams p_in, const float* __restrict__ in, float* __restrict__ out, int ns, int ng,
                     int stage_bytes, int tables_in_smem, const __grid_constant__ CUtensorMap tmap)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const unsigned full0 = smem_u32(smem_raw + (size_t)ns * stage_bytes);
    const unsigned empty0 = full0 + 8 * ns;
    const int tid = threadIdx.x;
    const int gt = CT / ng;
    KParams p = p_in;
    if constexpr (Geo::kStatic) {
        p.scratch_off = 0;              // registered schedules have no generic prime step: fold that path away
        if (MODE == dft::MODE_C2C_COLS) {
            Geo::apply(p.col);
            p.H = Geo::N;
        } else {
            Geo::apply(p.row);
            p.M = Geo::N;
            p.W = 2 * Geo::N;
            p.pad = (Geo::pow2 && Geo::N >= 32) ? 1 : 0;
            constexpr int PP = Geo::N + (Geo::N >> 4) + (Geo::N >> 8) + 1;
            p.P = (Geo::pow2 && Geo::N >= 32) ? PP : (MODE == dft::MODE_I1D ? Geo::N + 1 : Geo::N);
            if constexpr (!(Geo::pow2 && Geo::N >= 32)) {
                // same decision as the planner's (dft_planner.h: add_rows)
                Geo::template apply_orders<(MODE == dft::MODE_I1D ? Geo::N + 1 : Geo::N)>(p.row);
            }
        }
    }
    // Registered geometries are only dispatched when their tables fit (pipe_kernel_for_geo): the branch is then
    // compile-time true and the compiler knows that every table pointer is a shared-memory address (LDS, not LD).
    if (Geo::kStatic || tables_in_smem) {
        // twiddle / permutation tables move into shared memory once per (persistent) CTA
        unsigned char* tb = smem_raw + (size_t)ns * stage_bytes + 2 * ns * 8 + 64;
        auto stash = [&](const void* src, int bytes) -> void* {
            void* dst = tb;
            for (int i = tid * 4; i < bytes; i += (CT + PT) * 4) *(int*)(tb + i) = __ldg((const int*)((const unsigned char*)src + i));
            tb += (bytes + 15) / 16 * 16;
            return dst;
        };
        if (MODE == dft::MODE_C2C_COLS || MODE == dft::MODE_F2D || MODE == dft::MODE_I2D) {
            p.tw_col = (const float2*)stash(p_in.tw_col, p_in.col.n * 8);
            p.perm_col = (const int*)stash(p_in.perm_col, p_in.col.n * 4);
        }
        if (MODE != dft::MODE_C2C_COLS) {
            if constexpr (geo_two_rows<Geo>::value) {
                Geo::fill_table((float2*)tb, p_in.tw_row, tid, CT + PT);   // p.tw_row now means T (dft_twostep.cuh)
                p.tw_row = (const float2*)tb;
                tb += (p_in.row.n * 8 + 15) / 16 * 16;
            } else if (Geo::kStatic || p_in.tw_row_t) {           // (the planner always provides it for the row modes)
                p.tw_row_t = (const float2*)stash(p_in.tw_row_t, (p_in.row.n > 1 ? p_in.row.n - 1 : 1) * 8);   // (the plain table stays in global memory, unused)
            } else {
                p.tw_row = (const float2*)stash(p_in.tw_row, p_in.row.n * 8);
            }
            p.perm_row = (const int*)stash(p_in.perm_row, p_in.row.n * 4);
            // (bulk-store epilogue: the split twiddles stay in global memory, their 8 (M + 1) bytes go to the staging rows)
            if (!p_in.pack_bulk) p.tw_unt = (const float2*)stash(p_in.tw_unt, (p_in.M + 1) * 8);
        }
    }
    if (tid == 0) {
        for (int s = 0; s < ns; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full0 + 8 * s), "r"(PT) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(empty0 + 8 * s), "r"(gt / 32) : "memory");
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (tid >= CT) {
        const int pt = tid - CT;
        long long it = 0;
        for (long long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const int s = (int)(it % ns);
            if (it >= ns) gp_wait(empty0 + 8 * s, (unsigned)(((it / ns) - 1) & 1));
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            float2* buf = reinterpret_cast<float2*>(smem_raw + (size_t)s * stage_bytes);
            const unsigned bar = full0 + 8 * s;
            const unsigned bulk = produce_tile<MODE, PT>(p, tile, buf, in, pt, bar, &tmap);
            if (bulk) {
                if (pt != 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
            } else {
                asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
            }
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        return;
    }
    const int g = tid / gt;
    const GroupCtx ctx{tid - g * gt, gt, 1 + g,
                       p_in.pack_bulk ? reinterpret_cast<float2*>(smem_raw + p_in.pack_stage_off + (size_t)g * p_in.pack_stage_bytes) : nullptr};
    long long it = g;
    for (long long tile = blockIdx.x + (long long)g * gridDim.x; tile < p.ntiles; tile += (long long)ng * gridDim.x, it += ng) {
        const int s = (int)(it % ns);
        float2* buf = reinterpret_cast<float2*>(smem_raw + (size_t)s * stage_bytes);
        gp_wait(full0 + 8 * s, (unsigned)((it / ns) & 1));
        if constexpr (MODE == dft::MODE_F2D) dft::program_f2d<const GroupCtx, false>(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_I2D) dft::program_i2d<const GroupCtx, false>(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_F1D) {
            if constexpr (geo_two_rows<Geo>::value) Geo::run_f1d(p, tile, buf, out, ctx);
            else dft::program_f1d<const GroupCtx, false>(p, tile, buf, in, out, ctx);
        }
        if constexpr (MODE == dft::MODE_I1D) {
            if constexpr (geo_two_rows<Geo>::value) Geo::run_i1d(p, tile, buf, out, ctx);
            else dft::program_i1d<const GroupCtx, false>(p, tile, buf, in, out, ctx);
        }
        if constexpr (MODE == dft::MODE_C2C_COLS) {
            if constexpr (geo_two_step<Geo>::value) Geo::run_cols(p, tile, buf, out, ctx);
            else dft::program_cols<const GroupCtx, false>(p, tile, buf, in, out, ctx);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if ((tid & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty0 + 8 * s) : "memory");
    }
    // bulk stores out of the staging rows must have completed before the CTA's shared memory goes away
    if (p.pack_bulk && ctx.tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace

namespace dft {

// registered lengths: (N, radix schedule) must equal what the planner chose for that axis
#define DFT_GEO_ROWS(X)          \
    X(720, 16, 9, 5, 1)          \
    X(4096, 16, 16, 16, 1)       \
    X(1024, 4, 16, 16, 1)        \
    X(2048, 8, 16, 16, 1)
#define DFT_GEO_COLS(X) X(720, 10, 9, 8, 1)
// (N, planner schedule, R1A, R1B, R2A, R2B, TM, may be the first pass of a four-step FFT)
// (M, planner schedule, R1A, R1B, R2A, R2B): unskewed rows, plain output, tables in shared memory
#define DFT_GEO_ROWS_TWO(X) X(720, 16, 9, 5, 3, 8, 5, 6)
#define DFT_GEO_COLS_TWO(X)              \
    X(720, 10, 9, 8, 3, 8, 5, 6, 16, false)     \
    X(720, 10, 9, 8, 3, 8, 5, 6, 8, false)      \
    X(128, 16, 8, 1, 8, 1, 16, 1, 32, true)     \
    X(128, 16, 8, 1, 8, 1, 16, 1, 64, true)     \
    X(64, 8, 8, 1, 8, 1, 8, 1, 64, true)

pipe_fn pipe_kernel_for_geo(const KParams& k, bool tables_in_smem, int* compute_threads, int* producer_threads)
{
    *compute_threads = kPipeCompute;
    *producer_threads = kPipeProducers;
    if (!tables_in_smem) return nullptr;                 // the static kernels keep their tables in shared memory
    const int mode = k.mode;
    const dft::LineFFT& f = mode == dft::MODE_C2C_COLS ? k.col : k.row;
#define DFT_GEO_MATCH(N_, A, B, C, D)                                                                          \
    (f.n == N_ && f.nsteps == ((A > 1) + (B > 1) + (C > 1) + (D > 1)) && f.radix[0] == A && (B == 1 || f.radix[1] == B) && \
     (C == 1 || f.radix[2] == C) && (D == 1 || f.radix[3] == D))
#define DFT_GEO_TRY_ROWS(N_, A, B, C, D)                                                                      \
    if (DFT_GEO_MATCH(N_, A, B, C, D)) {                                                                      \
        if (mode == dft::MODE_F1D) return pipe_tile_kernel<dft::MODE_F1D, GeoLine<N_, A, B, C, D>>;           \
        if (mode == dft::MODE_I1D) return pipe_tile_kernel<dft::MODE_I1D, GeoLine<N_, A, B, C, D>>;           \
    }
#define DFT_GEO_TRY_COLS(N_, A, B, C, D)                                                                      \
    if (DFT_GEO_MATCH(N_, A, B, C, D) && mode == dft::MODE_C2C_COLS)                                          \
        return pipe_tile_kernel<dft::MODE_C2C_COLS, GeoLine<N_, A, B, C, D>>;
#define DFT_GEO_TRY_COLS_TWO(N_, A, B, C, R1A, R1B, R2A, R2B, TM_, FOUR_)                                            \
    if (DFT_GEO_MATCH(N_, A, B, C, 1) && mode == dft::MODE_C2C_COLS && k.G == TM_ && !(k.inverse && k.n1split > 1) && (FOUR_ || k.n1split <= 1) && tables_in_smem && \
        !getenv("TRT_DFT_NO_TWOSTEP")) {                                                                       \
        /* both steps' work items (R2*TM and R1*TM) fit 480 threads: 16 warps, 128 registers (see pipe_tile_kernel) */ \
        if constexpr ((R2A * R2B) * TM_ <= 480 && (R1A * R1B) * TM_ <= 480 && (R2A * R2B) * TM_ > 256) {       \
            if (getenv("TRT_DFT_480")) { /* A/B only: measured slower, one producer warp cannot feed the tile ring */ \
                *compute_threads = 480;                                                                        \
                *producer_threads = 32;                                                                        \
                return pipe_tile_kernel<dft::MODE_C2C_COLS, GeoTwo<N_, A, B, C, R1A, R1B, R2A, R2B, TM_, FOUR_>, 480, 32>; \
            }                                                                                                  \
        }                                                                                                      \
        /* four producer warps instead of two (20 warps still leave 96 registers) where the producers compute every row */ \
        /* address (index-reversed first pass of an inverse four-step FFT: cfg5 inverse 1.65 -> 1.57 ms; elsewhere +-1 %) */ \
        if (k.rev_in && !getenv("TRT_DFT_NO_P128")) {                                                          \
            *compute_threads = 512;                                                                            \
            *producer_threads = 128;                                                                           \
            return pipe_tile_kernel<dft::MODE_C2C_COLS, GeoTwo<N_, A, B, C, R1A, R1B, R2A, R2B, TM_, FOUR_>, 512, 128>; \
        }                                                                                                      \
        return pipe_tile_kernel<dft::MODE_C2C_COLS, GeoTwo<N_, A, B, C, R1A, R1B, R2A, R2B, TM_, FOUR_>>;    \
    }
    DFT_GEO_COLS_TWO(DFT_GEO_TRY_COLS_TWO)
#undef DFT_GEO_TRY_COLS_TWO
#define DFT_GEO_TRY_ROWS_TWO(N_, A, B, C, R1A, R1B, R2A, R2B)                                                     \
    if (DFT_GEO_MATCH(N_, A, B, C, 1) && k.pad == 0 && k.pack_n == 0 && tables_in_smem && !getenv("TRT_DFT_NO_TWOSTEP")) { \
        if (mode == dft::MODE_F1D) return pipe_tile_kernel<dft::MODE_F1D, GeoTwoRows<N_, A, B, C, R1A, R1B, R2A, R2B>>; \
        if (mode == dft::MODE_I1D) return pipe_tile_kernel<dft::MODE_I1D, GeoTwoRows<N_, A, B, C, R1A, R1B, R2A, R2B>>; \
    }
    DFT_GEO_ROWS_TWO(DFT_GEO_TRY_ROWS_TWO)
#undef DFT_GEO_TRY_ROWS_TWO
    DFT_GEO_ROWS(DFT_GEO_TRY_ROWS)
    DFT_GEO_COLS(DFT_GEO_TRY_COLS)
#undef DFT_GEO_TRY_ROWS
#undef DFT_GEO_TRY_COLS
#undef DFT_GEO_MATCH
    return nullptr;
}

pipe_fn pipe_kernel_for(int mode)
{
    switch (mode) {
        // (the fused 2-D programs gain only ~4 % under this shell -- measured -- so they stay on tile_kernel)
        case dft::MODE_F1D: return pipe_tile_kernel<dft::MODE_F1D>;
        case dft::MODE_I1D: return pipe_tile_kernel<dft::MODE_I1D>;
        case dft::MODE_C2C_COLS: return pipe_tile_kernel<dft::MODE_C2C_COLS>;
        default: return nullptr;
    }
}

}  // namespace dft
