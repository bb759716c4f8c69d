// dft_trunc.cu -- crop / zero-pad kernels behind the truncated-mode plans of shapes that have no fused truncated
// kernel: the full transform runs through the workspace and these move the kept modes
//     rows q in [0, kh) u [H - kh, H), columns k in [0, kw)
// between the full one-sided spectrum [img][H][MO] and the dense kept block [img][2*kh][kw].
#include <cuda_runtime.h>

#include "dft_launch.h"

namespace {

// one thread per kept element: full -> kept (forward) ; one thread per FULL element: kept -> full with zeros (inverse)
__global__ void __launch_bounds__(256) crop_modes_kernel(const float2* __restrict__ full, float2* __restrict__ kept, long long nimg, int H,
                                                        int MO, int kh, int kw)
{
    const long long per = 2LL * kh * kw, total = nimg * per;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long img = i / per;
        const int rem = (int)(i - img * per);
        const int qi = rem / kw, k = rem - qi * kw;
        const int q = qi < kh ? qi : qi + (H - 2 * kh);
        kept[i] = __ldg(full + (img * H + q) * MO + k);
    }
}

__global__ void __launch_bounds__(256) pad_modes_kernel(const float2* __restrict__ kept, float2* __restrict__ full, long long nimg, int H,
                                                       int MO, int kh, int kw)
{
    const long long per = (long long)H * MO, total = nimg * per;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long img = i / per;
        const int rem = (int)(i - img * per);
        const int q = rem / MO, k = rem - q * MO;
        const int qi = q < kh ? q : (q >= H - kh ? q - (H - 2 * kh) : -1);
        float2 v = float2{0.f, 0.f};
        if (qi >= 0 && k < kw) v = __ldg(kept + (img * 2 * kh + qi) * kw + k);
        full[i] = v;
    }
}

int grid_for(long long total)
{
    long long b = (total + 255) / 256;
    if (b > 148LL * 32) b = 148LL * 32;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace

namespace dft {

void launch_crop_modes(const void* full, void* kept, long long nimg, int H, int MO, int kh, int kw, void* stream)
{
    crop_modes_kernel<<<grid_for(nimg * 2 * kh * kw), 256, 0, (cudaStream_t)stream>>>((const float2*)full, (float2*)kept, nimg, H, MO, kh, kw);
}

void launch_pad_modes(const void* kept, void* full, long long nimg, int H, int MO, int kh, int kw, void* stream)
{
    pad_modes_kernel<<<grid_for(nimg * H * MO), 256, 0, (cudaStream_t)stream>>>((const float2*)kept, (float2*)full, nimg, H, MO, kh, kw);
}

}  // namespace dft
