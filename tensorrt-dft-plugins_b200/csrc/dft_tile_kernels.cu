// dft_tile_kernels.cu -- the multi-CTA-per-SM shell of the generic engine and the naive fallback kernels.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "dft_launch.h"
#include "dft_planner.h"
#include "dft_programs.cuh"

namespace {

using dft::KParams;

__device__ __forceinline__ unsigned smem_u32(const void* p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}

struct CtaCtx {
    int tid, nthr;
    unsigned bar;      // shared-memory address of the mbarrier
    unsigned phase;    // parity of the next completion
    bool bulk;         // this mode may use cp.async.bulk

    __host__ __device__ __forceinline__ void sync()
    {
#if defined(__CUDA_ARCH__)
        __syncthreads();
#endif
    }

    // bulk stores out of staging rows exist in the pipelined shell only (KParams::pack_bulk is never set here)
    __host__ __device__ __forceinline__ float2* stage_out() const { return nullptr; }
    __host__ __device__ __forceinline__ void fence_async_shared() const {}
    __host__ __device__ __forceinline__ void bulk_store(void*, const void*, unsigned) const {}
    __host__ __device__ __forceinline__ void commit_bulk_stores() const {}
    __host__ __device__ __forceinline__ void wait_bulk_stores_read() const {}

    // Whole-tile global->shared copy by the TMA unit (1-D bulk copy).  All threads call it with
    // identical arguments; returns false (nothing done) when the copy is not 16-byte clean.
    __host__ __device__ __forceinline__ bool bulk_load(void* dst, const void* src, long long bytes)
    {
#if defined(__CUDA_ARCH__)
        if (!bulk || bytes <= 0 || (bytes & 15) || ((uintptr_t)src & 15) || bytes >= (1 << 20)) return false;
        if (tid == 0) {
            // order the previous tile's generic-proxy accesses before the async-proxy write
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)bytes)
                         : "memory");
            asm volatile(
                "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                    smem_u32(dst)),
                "l"(src), "r"((unsigned)bytes), "r"(bar)
                : "memory");
        }
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "DFT_WAIT:\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
            "@p bra DFT_DONE;\n"
            "bra DFT_WAIT;\n"
            "DFT_DONE:\n"
            "}\n" ::"r"(bar),
            "r"(phase)
            : "memory");
        phase ^= 1u;
        return true;
#else
        (void)dst; (void)src; (void)bytes;
        return false;
#endif
    }
};

template <int MODE, int NT>
__global__ void __launch_bounds__(NT, NT == 512 ? 2 : 3) tile_kernel(const __grid_constant__ KParams p, const float* __restrict__ in,
                                                              float* __restrict__ out)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* buf = reinterpret_cast<float2*>(smem_raw);
    CtaCtx ctx;
    ctx.tid = threadIdx.x;
    ctx.nthr = blockDim.x;
    ctx.phase = 0;
    ctx.bulk = (MODE == dft::MODE_F2D || MODE == dft::MODE_I2D || MODE == dft::MODE_F1D || MODE == dft::MODE_I1D);
    ctx.bar = smem_u32(smem_raw + p.smem_bytes - dft::kBarrierBytes);
    if (ctx.bulk) {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ctx.bar) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        __syncthreads();
    }
    for (long long tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        if constexpr (MODE == dft::MODE_F2D) dft::program_f2d(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_I2D) dft::program_i2d(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_F1D) dft::program_f1d(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_I1D) dft::program_i1d(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_F1D_ODD) dft::program_f1d_odd(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_I1D_ODD) dft::program_i1d_odd(p, tile, buf, in, out, ctx);
        if constexpr (MODE == dft::MODE_C2C_COLS) dft::program_cols(p, tile, buf, in, out, ctx);
    }
}

template <int MODE>
__global__ void __launch_bounds__(256) naive_kernel(const __grid_constant__ KParams p, const float* __restrict__ in,
                                                    float* __restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < p.nunits; i += stride) {
        if constexpr (MODE == dft::MODE_NAIVE_ROWS_R2C) dft::naive_rows_r2c_elem(p, i, in, out);
        if constexpr (MODE == dft::MODE_NAIVE_ROWS_C2R) dft::naive_rows_c2r_elem(p, i, in, out);
        if constexpr (MODE == dft::MODE_NAIVE_COLS) dft::naive_cols_elem(p, i, in, out);
    }
}

// ---- a handful of short rows (BASELINE config 1, [4,32,64]): launch-latency bound -------------------------------------
// One thread per row: the row's M = W/2 packed complex points live in registers from the global loads to the global
// stores -- no shared memory, no barrier, no table, a single dependent chain load -> butterflies -> store.  The
// accesses are not coalesced (consecutive threads are a whole row apart), which is irrelevant for the few KB this
// kernel is planned for and is why the planner only picks it for small batches.
template <int M, bool INV>
__global__ void __launch_bounds__(64) small_rows_kernel(const __grid_constant__ KParams p, const float* __restrict__ in, float* __restrict__ out)
{
    using namespace dft;
    constexpr int W = 2 * M;
    const long long r = blockIdx.x * 64LL + threadIdx.x;
    if (r >= p.nunits) return;
    float2 z[M];
    if (!INV) {
        const float* src = in + r * W;
        if ((((uintptr_t)src) & 15) == 0 && M % 2 == 0) {
            static_for<0, M / 2>([&](auto ic) {
                constexpr int i = decltype(ic)::value;
                const float4 q = __ldg(reinterpret_cast<const float4*>(src) + i);
                z[2 * i] = float2{q.x, q.y};
                z[2 * i + 1] = float2{q.z, q.w};
            });
        } else {
            static_for<0, M>([&](auto ic) {
                constexpr int i = decltype(ic)::value;
                z[i] = float2{__ldg(src + 2 * i), __ldg(src + 2 * i + 1)};
            });
        }
        dft_reg<M>(z);
        float2* o = reinterpret_cast<float2*>(out) + r * (M + 1);
        o[0] = float2{z[0].x + z[0].y, 0.f};
        o[M] = float2{z[0].x - z[0].y, 0.f};
        static_for<1, M / 2 + 1>([&](auto kc) {
            constexpr int k = decltype(kc)::value;
            constexpr float wx = float(cx_cos2pi(k, W)), wy = float(-cx_sin2pi(k, W));      // w_W^k
            const float2 a = z[k], b = z[M - k];
            const float sx = a.x + b.x, sy = a.y - b.y, dx = a.x - b.x, dy = a.y + b.y;
            const float tx = 0.5f * (wx * dx - wy * dy), ty = 0.5f * (wx * dy + wy * dx);
            o[k] = float2{0.5f * sx + ty, 0.5f * sy - tx};
            if constexpr (k != M - k) o[M - k] = float2{0.5f * sx - ty, -0.5f * sy - tx};
        });
    } else {
        const float2* src = reinterpret_cast<const float2*>(in) + r * (M + 1);
        {
            const float a = __ldg(src).x, b = __ldg(src + M).x;       // the imaginary parts of DC / Nyquist are ignored
            z[0] = float2{a + b, a - b};
        }
        static_for<1, M / 2 + 1>([&](auto kc) {
            constexpr int k = decltype(kc)::value;
            constexpr float wx = float(cx_cos2pi(k, W)), wy = float(-cx_sin2pi(k, W));
            const float2 a = __ldg(src + k), b = __ldg(src + (M - k));
            const float sx = a.x + b.x, sy = a.y - b.y, dx = a.x - b.x, dy = a.y + b.y;
            const float ux = wx * dx + wy * dy, uy = wx * dy - wy * dx;
            z[k] = float2{sx - uy, sy + ux};
            if constexpr (k != M - k) z[M - k] = float2{sx + uy, -sy + ux};
        });
        dft_reg<M>(z);
        const float sc = p.scale;
        float2* o = reinterpret_cast<float2*>(out) + r * M;          // real row as pairs (x[2n], x[2n+1]) = y[(M - n) mod M]
        static_for<0, M>([&](auto nc) {
            constexpr int n = decltype(nc)::value;
            constexpr int src_i = (M - n) % M;
            o[n] = float2{z[src_i].x * sc, z[src_i].y * sc};
        });
    }
}

}  // namespace

namespace dft {

kernel_fn small_rows_kernel_for(int mode, int M)
{
    const bool inv = mode == MODE_SMALL_ROWS_C2R;
    switch (M) {
#define DFT_SMALL(M_) \
    case M_: return inv ? (kernel_fn)small_rows_kernel<M_, true> : (kernel_fn)small_rows_kernel<M_, false>;
        DFT_SMALL(2)
        DFT_SMALL(4)
        DFT_SMALL(8)
        DFT_SMALL(16)
        DFT_SMALL(32)
#undef DFT_SMALL
        default: return nullptr;
    }
}

kernel_fn tile_kernel_for(int mode, int threads)
{
    switch (mode) {
        case dft::MODE_F2D: return tile_kernel<dft::MODE_F2D, 256>;
        case dft::MODE_I2D: return tile_kernel<dft::MODE_I2D, 256>;
        case dft::MODE_F1D: return tile_kernel<dft::MODE_F1D, 256>;
        case dft::MODE_I1D: return tile_kernel<dft::MODE_I1D, 256>;
        case dft::MODE_F1D_ODD: return tile_kernel<dft::MODE_F1D_ODD, 256>;
        case dft::MODE_I1D_ODD: return tile_kernel<dft::MODE_I1D_ODD, 256>;
        case dft::MODE_C2C_COLS:
            return threads == 512 ? tile_kernel<dft::MODE_C2C_COLS, 512> : tile_kernel<dft::MODE_C2C_COLS, 256>;
        case dft::MODE_NAIVE_ROWS_R2C: return naive_kernel<dft::MODE_NAIVE_ROWS_R2C>;
        case dft::MODE_NAIVE_ROWS_C2R: return naive_kernel<dft::MODE_NAIVE_ROWS_C2R>;
        case dft::MODE_NAIVE_COLS: return naive_kernel<dft::MODE_NAIVE_COLS>;
        default: return nullptr;
    }
}

// ---- shape-specialised fused 2-D kernels (dft_fused2d.cuh) ------------------------------------
}  // namespace dft
