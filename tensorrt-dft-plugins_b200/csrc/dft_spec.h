// dft_spec.h -- registry of shape-specialised fused 2-D kernels (instantiated in dft_spec.cu so that they
// compile in parallel with the generic engine in dft_kernels.cu).
#pragma once
#include <vector_types.h>

namespace dft {

typedef void (*spec_fn)(const float*, float*, long long, const float2*, float);
// truncated-mode variants: (in, out, images, tables, scale, keep_h, keep_w); the forward ignores scale
typedef void (*trunc_fn)(const float*, float*, long long, const float2*, float, int, int);

struct SpecEntry {
    int H, W, RA, RB, CA, CB, NT;
    int smem_f, smem_i;
    spec_fn fwd, inv;
    bool pipelined;      // dft_pipe2d.cuh: one persistent CTA per SM, needs 16-byte aligned in/out
    int NT_inv = 0;      // threads of the inverse kernel when they differ from NT
    trunc_fn trunc_fwd = nullptr;   // same CTA size and shared memory as fwd / inv
    trunc_fn trunc_inv = nullptr;
    // 16-bit tensors on both sides, fp32 arithmetic ([0] fp16, [1] bf16; the pointers are passed as float*): same CTA
    // size and shared memory again
    spec_fn fwd16[2] = {nullptr, nullptr};
    spec_fn inv16[2] = {nullptr, nullptr};
};

// nullptr when (H, W) has no specialised kernel
const SpecEntry* find_spec(int H, int W);

}  // namespace dft
