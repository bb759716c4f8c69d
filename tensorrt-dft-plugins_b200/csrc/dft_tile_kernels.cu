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


}  // namespace

namespace dft {

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
