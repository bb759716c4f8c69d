// dft_spec.cu -- instantiations of the shape-specialised fused rfft2 / irfft2 kernels.
#include <stdlib.h>

#include "dft_fused2d.cuh"
#include "dft_pipe2d.cuh"
#include "dft_reg2d.cuh"
#include "dft_spec.h"

namespace dft {
namespace {

template <class C>
SpecEntry make_spec()
{
    SpecEntry e{C::H, C::W, C::RA, C::RB, C::CA, C::CB, C::NT, C::SMEM_F, C::SMEM_I,
                fused_rfft2_kernel<C, float>, fused_irfft2_kernel<C, float>, false};
    e.trunc_fwd = fused_rfft2_kernel<C, float, int, int>;
    e.trunc_inv = fused_irfft2_kernel<C, float, int, int>;
    e.fwd16[0] = fused_rfft2_kernel<C, __half>;
    e.inv16[0] = fused_irfft2_kernel<C, __half>;
    e.fwd16[1] = fused_rfft2_kernel<C, __nv_bfloat16>;
    e.inv16[1] = fused_irfft2_kernel<C, __nv_bfloat16>;
    return e;
}

template <class CF, class CI>
SpecEntry make_pipe()
{
    static_assert(CF::H == CI::H && CF::W == CI::W && CF::RA == CI::RA && CF::RB == CI::RB && CF::CA == CI::CA && CF::CB == CI::CB, "");
    SpecEntry e{CF::H, CF::W, CF::RA, CF::RB, CF::CA, CF::CB, CF::THREADS, CF::SMEM, CI::SMEM,
                pipe_rfft2_kernel<CF>, pipe_irfft2_kernel<CI>, true};
    e.NT_inv = CI::THREADS;
    return e;
}

// both kernels keep whole lines in registers (dft_reg2d.cuh); they need no tables
template <class C>
SpecEntry make_reg()
{
    SpecEntry e{C::H, C::W, C::RA, C::RB, C::CA, C::CB, C::THREADS, C::SMEM_F, C::SMEM_I,
                reg_rfft2_kernel<C>, reg_irfft2_kernel<C>, true};
    e.trunc_fwd = reg_rfft2_trunc_kernel<C>;
    e.trunc_inv = reg_irfft2_trunc_kernel<C>;
    e.fwd16[0] = reg_rfft2_io16_kernel<C, __half>;
    e.inv16[0] = reg_irfft2_io16_kernel<C, __half>;
    e.fwd16[1] = reg_rfft2_io16_kernel<C, __nv_bfloat16>;
    e.inv16[1] = reg_irfft2_io16_kernel<C, __nv_bfloat16>;
    return e;
}

}  // namespace

const SpecEntry* find_spec(int H, int W)
{
    static const SpecEntry reg_table[] = {
        //               H    W   RA RB CA CB groups stages I/O-warps ping-pong
        make_reg<RegCfg<90, 180, 9, 5, 9, 5, 2, 3, true, 1>>(),         // FourCastNet AFNO (BASELINE cfg3)
    };
    if (!getenv("TRT_DFT_NO_REG2D"))
        for (const auto& e : reg_table)
            if (e.H == H && e.W == W) return &e;
    static const SpecEntry table[] = {
        //                H    W   RA  RB  CA  CB  group threads, groups, stages
        make_pipe<PipeCfg<90, 180, 9, 10, 10, 9, 320, 2, 3>,      // forward: 2 x 320 threads
                  PipeCfg<90, 180, 9, 10, 10, 9, 480, 2, 3>>(),    // inverse: 2 x 480 (one-round merge)   FourCastNet AFNO (BASELINE cfg3)
        //                 H    W   RA  RB  CA  CB   NT  min CTAs/SM
        make_spec<FusedCfg<128, 128, 8, 8, 16, 8, 256, 3>>(),     // FNO (BASELINE cfg2)
        make_spec<FusedCfg<64, 64, 8, 4, 8, 8, 256, 4>>(),
        make_spec<FusedCfg<64, 128, 8, 8, 8, 8, 256, 4>>(),       // rectangular FNO grids
        make_spec<FusedCfg<128, 64, 8, 4, 16, 8, 256, 3>>(),
    };
    const bool no_pipe = getenv("TRT_DFT_NO_PIPE") != nullptr;   // benchmarking override
    for (const auto& e : table)
        if (e.H == H && e.W == W && !(no_pipe && e.pipelined)) return &e;
    return nullptr;
}

}  // namespace dft
