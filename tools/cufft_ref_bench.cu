// tools/cufft_ref_bench.cu -- COMPARATOR ONLY (never linked into libtrt_dft_plugins.so).
//
// Issues, without TensorRT, the call sequence the reference plugin makes:
//   configurePlugin: cufftCreate, cufftSetAutoAllocation(0), cufftXtMakePlanMany(rank, n, NULL,1,0,
//                    in_t, NULL,1,0, out_t, batch, &ws, CUDA_C_32F)   (dft_plugins.cpp:161-176)
//   enqueue:         cufftSetStream, cufftSetWorkArea, cufftXtExec     (dft_plugins.cpp:185-195)
//   Irfft enqueue:   + cublasSetStream, cublasScalEx(1/N)              (dft_plugins.cpp:454-468)
// and times it with CUDA events, so "the reference's cuFFT path" has a number on this box.
//
//   nvcc -O2 -o tools/cufft_ref_bench tools/cufft_ref_bench.cu -lcufft -lcublas
//   tools/cufft_ref_bench <iters> <batch> <d0> [d1 [d2]]
#include <cublas_v2.h>
#include <cuda_runtime.h>
#include <cufftXt.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x) do { auto e_ = (x); if (e_ != 0) { fprintf(stderr, "%s failed: %d (line %d)\n", #x, (int)e_, __LINE__); exit(2); } } while (0)

// Parity mode: the same call sequence on a caller-supplied input, results written to files, so that the kernels of
// libtrt_dft_plugins.so can be compared with what the reference's enqueue would have produced on this box.
//   tools/cufft_ref_bench check <x.f32> <fwd_out.f32> <inv_out.f32> batch d0 [d1 [d2]]
// x.f32: batch*prod(d) floats.  fwd_out = cufftXtExec(R2C)(x); inv_out = scal(1/N) * cufftXtExec(C2R)(fwd_out).
static int check_mode(int argc, char** argv)
{
    if (argc < 7) { fprintf(stderr, "usage: %s check x.f32 fwd.f32 inv.f32 batch d0 [d1 [d2]]\n", argv[0]); return 1; }
    const long long batch = atoll(argv[5]);
    const int rank = argc - 6;
    long long n[3] = {1, 1, 1};
    double total = 1.0;
    for (int i = 0; i < rank; ++i) { n[i] = atoll(argv[6 + i]); total *= (double)n[i]; }
    long long lead = 1;
    for (int i = 0; i + 1 < rank; ++i) lead *= n[i];
    const long long W = n[rank - 1], Mo = W / 2 + 1;
    const size_t real_bytes = (size_t)batch * lead * W * 4, cplx_bytes = (size_t)batch * lead * Mo * 8;
    float* hx = (float*)malloc(real_bytes);
    float* hy = (float*)malloc(cplx_bytes);
    FILE* fi = fopen(argv[2], "rb");
    if (!fi || fread(hx, 1, real_bytes, fi) != real_bytes) { fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
    fclose(fi);
    float *x, *y, *xr;
    CK(cudaMalloc(&x, real_bytes)); CK(cudaMalloc(&y, cplx_bytes)); CK(cudaMalloc(&xr, real_bytes));
    CK(cudaMemcpy(x, hx, real_bytes, cudaMemcpyHostToDevice));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cufftHandle hf, hi; size_t wsf = 0, wsi = 0;
    CK(cufftCreate(&hf)); CK(cufftSetAutoAllocation(hf, 0));
    CK(cufftXtMakePlanMany(hf, rank, n, nullptr, 1, 0, CUDA_R_32F, nullptr, 1, 0, CUDA_C_32F, batch, &wsf, CUDA_C_32F));
    CK(cufftCreate(&hi)); CK(cufftSetAutoAllocation(hi, 0));
    CK(cufftXtMakePlanMany(hi, rank, n, nullptr, 1, 0, CUDA_C_32F, nullptr, 1, 0, CUDA_R_32F, batch, &wsi, CUDA_C_32F));
    void *wf = nullptr, *wi = nullptr;
    if (wsf) CK(cudaMalloc(&wf, wsf));
    if (wsi) CK(cudaMalloc(&wi, wsi));
    cublasHandle_t cb; CK(cublasCreate(&cb));
    const float scale = 1.0f / (float)total;
    CK(cufftSetStream(hf, st)); CK(cufftSetWorkArea(hf, wf)); CK(cufftXtExec(hf, x, y, CUFFT_FORWARD));
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(hy, y, cplx_bytes, cudaMemcpyDeviceToHost));      // before the C2R, which may overwrite its input
    CK(cufftSetStream(hi, st)); CK(cufftSetWorkArea(hi, wi)); CK(cufftXtExec(hi, y, xr, CUFFT_INVERSE));
    CK(cublasSetStream(cb, st));
    CK(cublasScalEx(cb, (int)(batch * total), &scale, CUDA_R_32F, xr, CUDA_R_32F, 1, CUDA_R_32F));
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(hx, xr, real_bytes, cudaMemcpyDeviceToHost));
    FILE* fo = fopen(argv[3], "wb");
    if (!fo || fwrite(hy, 1, cplx_bytes, fo) != cplx_bytes) { fprintf(stderr, "cannot write %s\n", argv[3]); return 2; }
    fclose(fo);
    fo = fopen(argv[4], "wb");
    if (!fo || fwrite(hx, 1, real_bytes, fo) != real_bytes) { fprintf(stderr, "cannot write %s\n", argv[4]); return 2; }
    fclose(fo);
    printf("{\"impl\": \"cufft_reference_sequence\", \"mode\": \"check\", \"batch\": %lld}\n", batch);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc >= 2 && !strcmp(argv[1], "check")) return check_mode(argc, argv);
    if (argc < 4) { fprintf(stderr, "usage: %s iters batch d0 [d1 [d2]]\n", argv[0]); return 1; }
    const int iters = atoi(argv[1]);
    const long long batch = atoll(argv[2]);
    const int rank = argc - 3;
    long long n[3] = {1, 1, 1};
    double total = 1.0;
    for (int i = 0; i < rank; ++i) { n[i] = atoll(argv[3 + i]); total *= (double)n[i]; }
    long long lead = 1;
    for (int i = 0; i + 1 < rank; ++i) lead *= n[i];
    const long long W = n[rank - 1], Mo = W / 2 + 1;
    const size_t real_bytes = (size_t)batch * lead * W * 4, cplx_bytes = (size_t)batch * lead * Mo * 8;
    float *x, *y, *xr;
    CK(cudaMalloc(&x, real_bytes)); CK(cudaMalloc(&y, cplx_bytes)); CK(cudaMalloc(&xr, real_bytes));
    CK(cudaMemset(x, 0, real_bytes));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cufftHandle hf, hi; size_t wsf = 0, wsi = 0;
    CK(cufftCreate(&hf)); CK(cufftSetAutoAllocation(hf, 0));
    CK(cufftXtMakePlanMany(hf, rank, n, nullptr, 1, 0, CUDA_R_32F, nullptr, 1, 0, CUDA_C_32F, batch, &wsf, CUDA_C_32F));
    CK(cufftCreate(&hi)); CK(cufftSetAutoAllocation(hi, 0));
    CK(cufftXtMakePlanMany(hi, rank, n, nullptr, 1, 0, CUDA_C_32F, nullptr, 1, 0, CUDA_R_32F, batch, &wsi, CUDA_C_32F));
    void *wf = nullptr, *wi = nullptr;
    if (wsf) CK(cudaMalloc(&wf, wsf));
    if (wsi) CK(cudaMalloc(&wi, wsi));
    cublasHandle_t cb; CK(cublasCreate(&cb));
    const float scale = 1.0f / (float)total;
    auto fwd = [&]() { CK(cufftSetStream(hf, st)); CK(cufftSetWorkArea(hf, wf)); CK(cufftXtExec(hf, x, y, CUFFT_FORWARD)); };
    auto inv = [&]() {
        CK(cufftSetStream(hi, st)); CK(cufftSetWorkArea(hi, wi)); CK(cufftXtExec(hi, y, xr, CUFFT_INVERSE));
        CK(cublasSetStream(cb, st));
        CK(cublasScalEx(cb, (int)(batch * total), &scale, CUDA_R_32F, xr, CUDA_R_32F, 1, CUDA_R_32F));
    };
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float ms_f = 0, ms_i = 0, ms_p = 0;
    for (int i = 0; i < 5; ++i) { fwd(); inv(); }
    CK(cudaStreamSynchronize(st));
    CK(cudaEventRecord(a, st)); for (int i = 0; i < iters; ++i) fwd(); CK(cudaEventRecord(b, st));
    CK(cudaStreamSynchronize(st)); CK(cudaEventElapsedTime(&ms_f, a, b));
    CK(cudaEventRecord(a, st)); for (int i = 0; i < iters; ++i) inv(); CK(cudaEventRecord(b, st));
    CK(cudaStreamSynchronize(st)); CK(cudaEventElapsedTime(&ms_i, a, b));
    CK(cudaEventRecord(a, st)); for (int i = 0; i < iters; ++i) { fwd(); inv(); } CK(cudaEventRecord(b, st));
    CK(cudaStreamSynchronize(st)); CK(cudaEventElapsedTime(&ms_p, a, b));
    const double algo = (double)real_bytes + (double)cplx_bytes;
    printf("{\"impl\": \"cufft_reference_sequence\", \"rank\": %d, \"dims\": [%lld,%lld,%lld], \"batch\": %lld, "
           "\"fwd_ms\": %.5f, \"inv_ms\": %.5f, \"pair_ms\": %.5f, \"fwd_GBs\": %.1f, \"inv_GBs\": %.1f, \"pair_GBs\": %.1f, "
           "\"ws_fwd\": %zu, \"ws_inv\": %zu}\n",
           rank, n[0], n[1], n[2], batch, ms_f / iters, ms_i / iters, ms_p / iters, algo / (ms_f / iters * 1e-3) / 1e9,
           algo / (ms_i / iters * 1e-3) / 1e9, 2 * algo / (ms_p / iters * 1e-3) / 1e9, wsf, wsi);
    return 0;
}
