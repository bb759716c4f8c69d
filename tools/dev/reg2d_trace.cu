// Debug harness: phase timeline (clock64) of CTA 0 of the register-resident rfft2 kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -DDFT_REG2D_TRACE -Itensorrt-dft-plugins_b200/csrc tools/dev/reg2d_trace.cu -o gpurun_out/reg2d_trace
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "dft_reg2d.cuh"
using namespace dft;
#ifndef CFG_NG
#define CFG_NG 2
#endif
#ifndef CFG_IOW
#define CFG_IOW false
#endif
#ifndef CFG_PP
#define CFG_PP 0
#endif
using C = RegCfg<90, 180, 9, 5, 9, 5, CFG_NG, 3, CFG_IOW, CFG_PP>;
int main()
{
    const long long nimg = 6144;
    float *in, *out;
    long long* tr;
    cudaMalloc(&in, nimg * C::TILE_I);
    cudaMalloc(&out, nimg * C::TILE_I);
    cudaMemset(in, 0, nimg * C::TILE_I);
    const int slots = 64;
    cudaMalloc(&tr, slots * 8 * sizeof(long long));
    cudaMemset(tr, 0, slots * 8 * sizeof(long long));
    cudaMemcpyToSymbol(reg2d::g_trace, &tr, sizeof(tr));
    cudaFuncSetAttribute(reg_rfft2_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_F);
    for (int rep = 0; rep < 3; ++rep) reg_rfft2_kernel<C><<<148, C::THREADS, C::SMEM_F>>>(in, out, nimg, nullptr, 1.f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 20; ++rep) reg_rfft2_kernel<C><<<148, C::THREADS, C::SMEM_F>>>(in, out, nimg, nullptr, 1.f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("err %s, forward %.4f ms per launch\n", cudaGetErrorString(cudaGetLastError()), ms / 20);
    cudaFuncSetAttribute(reg_irfft2_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_I);
    for (int rep = 0; rep < 3; ++rep) reg_irfft2_kernel<C><<<148, C::THREADS, C::SMEM_I>>>(out, in, nimg, nullptr, 1.f);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 20; ++rep) reg_irfft2_kernel<C><<<148, C::THREADS, C::SMEM_I>>>(out, in, nimg, nullptr, 1.f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    cudaEventElapsedTime(&ms, e0, e1);
    printf("err %s, inverse %.4f ms per launch\n", cudaGetErrorString(cudaGetLastError()), ms / 20);
    std::vector<long long> h(slots * 8);
    cudaMemcpy(h.data(), tr, h.size() * 8, cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    printf("img grp | wait_full  loads  bar1  fft+split+sts  bar2  colloads  fft+stg | start end\n");
    for (int i = 0; i < (getenv("TRACE") ? 42 : 0); ++i) {
        long long* r = &h[i * 8];
        if (!r[7]) continue;
        printf("%3d  %d  | %6lld %6lld %6lld %6lld %6lld %6lld %6lld | %7lld %7lld\n", i, i % CFG_NG, r[1] - r[0], r[2] - r[1], r[3] - r[2], r[4] - r[3],
               r[5] - r[4], r[6] - r[5], r[7] - r[6], r[0] - t0, r[7] - t0);
    }
    return 0;
}
